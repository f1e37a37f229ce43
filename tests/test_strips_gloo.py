"""Host-side logic of the multi-GPU path on CPU: the strip planner (various_image_processings_b200/strips.py) and a
world-size-2 `gloo` run in which every rank filters only the rows it holds (own rows + halo) and the gathered result
must equal the whole-image result bit for bit.  The per-strip compute here is the ORACLE (this is a test of the
decomposition, not of the kernels; the CUDA strip entry points are tested in test_gpu_parity.py)."""
import os
import socket
import sys

import numpy as np
import pytest

import _oracle as O
from various_image_processings_b200 import halo_rows, plan_strips


def test_plan_covers_every_row_once():
    for h in (1, 7, 64, 131, 2160, 4320):
        for n in (1, 2, 3, 4, 8):
            for halo in (0, 4, 15, 17, 51):
                plans = plan_strips(h, n, halo)
                assert plans[0].y0 == 0 and plans[-1].y1 == h
                for a, b in zip(plans, plans[1:]):
                    assert a.y1 == b.y0
                for p in plans:
                    if p.y1 > p.y0:
                        assert p.s0 == max(0, p.y0 - halo) and p.s1 == min(h, p.y1 + halo)
                # balanced: strip heights differ by at most one row
                sizes = [p.y1 - p.y0 for p in plans]
                assert max(sizes) - min(sizes) <= 1


def test_halo_rows():
    assert halo_rows("bf", 31) == 15 and halo_rows("abf", 15) == 7 and halo_rows("gradient", 3) == 1
    assert halo_rows("btf", 9) == 17                       # (k-1) + k/2 + k/2 + 1
    assert halo_rows("btf", 9, nitr=3, exchange_between_iterations=False) == 51
    with pytest.raises(ValueError):
        plan_strips(0, 2, 1)


def _worker(rank, world, port, kind, k, nitr, q):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        w, h = 96, 90
        img = O.smooth_image(w, h, seed=11)
        halo = halo_rows(kind, k, nitr, exchange_between_iterations=False)
        p = plan_strips(h, world, halo)[rank]
        crop = np.ascontiguousarray(img[p.s0:p.s1])        # the rows this rank holds
        if kind == "bf":
            out = O.o_bf(crop, k)
        elif kind == "abf":
            out = O.o_abf(crop, k)
        else:
            out = O.o_btf(crop, k, nitr)
        own = torch.from_numpy(np.ascontiguousarray(out[p.y0 - p.s0:p.y1 - p.s0]))
        rows = [plan_strips(h, world, halo)[r] for r in range(world)]
        bufs = [torch.empty((r.y1 - r.y0, w, 3), dtype=torch.uint8) for r in rows]
        dist.all_gather(bufs, own) if len({b.shape for b in bufs}) == 1 else None
        if len({b.shape for b in bufs}) != 1:               # ragged strips: gather by broadcast
            for r in range(world):
                if r == rank:
                    bufs[r].copy_(own)
                dist.broadcast(bufs[r], src=r)
        got = torch.cat(bufs).numpy()
        want = {"bf": lambda: O.o_bf(img, k), "abf": lambda: O.o_abf(img, k), "btf": lambda: O.o_btf(img, k, nitr)}[kind]()
        q.put((rank, bool(np.array_equal(got, want))))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kind,k,nitr", [("bf", 11, 0), ("abf", 15, 0), ("btf", 5, 2)])
def test_world_size_2_gloo_strips_equal_whole_image(kind, k, nitr):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, kind, k, nitr, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(results) == [(0, True), (1, True)]
