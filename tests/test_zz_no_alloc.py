"""GPU test of vip_integral_u8c3_noalloc, the integral image for hosts that must not allocate device memory.

Who needs it: code that runs inside somebody else's memory budget -- a filter plug-in in an application that has already
handed the whole GPU to its own allocator, a frame loop that is forbidden to touch the driver's allocator after
start-up, a process whose memory is tracked by an outer framework.  The regular entry point (vip_integral_u8c3) takes a
few MB from a CUDA memory pool on every call; the variant keeps those aggregates in a fixed 20 MiB array inside the
library image and allocates nothing itself.  The caller still provides the image and the table (here: torch tensors).

What is checked, through the C ABI (ctypes), on cuda:0:
  * the table is identical, element for element, to the one vip_integral_u8c3 produces (integer sums: the bar in this
    repo is exact equality, see tests/test_gpu_parity.py) -- on TWO calls in a row, the second on a smaller image with
    a different radius, so anything the first call left in the shared array would show up;
  * a third call repeats the first image and must reproduce its table (the array keeps old contents between calls);
  * sizes whose aggregates exceed the array are refused with VIP_E_SIZE and nothing is written.

The file name sorts last on purpose: pytest -x stops at the first failure, and a failure here must not hide the parity
tests.  It passes on its own in a fresh process (first use of the library on the device) and after the other tests.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def L():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from various_image_processings_b200 import capi
    torch.cuda.set_device(0)
    return capi.lib()


def _image(w, h, seed):
    rng = np.random.default_rng(seed)
    return torch.from_numpy(rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)).cuda()


def _table(L, fn, src, w, h, r, fill):
    """Enqueue one build of the (h+2r+1) x (w+2r+1) x 3 table with entry point `fn`; returns the device tensor."""
    n = L.vip_integral_table_elems(w, h, r)
    assert n == (w + 2 * r + 1) * (h + 2 * r + 1) * 3
    out = torch.full((n,), fill, dtype=torch.int32, device="cuda")
    assert getattr(L, fn)(src.data_ptr(), out.data_ptr(), w, h, r, None) == 0
    return out


def _tables(L, src, w, h, r):
    """(table from the pool-backed entry point, table from the no-alloc entry point), both as int32 numpy arrays."""
    want = _table(L, "vip_integral_u8c3", src, w, h, r, -1)
    got = _table(L, "vip_integral_u8c3_noalloc", src, w, h, r, -2)
    torch.cuda.synchronize()
    return want.cpu().numpy(), got.cpu().numpy()


def test_noalloc_matches_pool_backed_entry_point_twice_in_a_row(L):
    # a: several 256-column chunks and several 32-row bands, neither size a multiple of the tile sizes
    # b: smaller frame, other radius, a single chunk -> a different carve-up of the same array
    a, b = _image(1337, 301, seed=11), _image(200, 97, seed=12)
    want_a = _table(L, "vip_integral_u8c3", a, 1337, 301, 7, -1)
    want_b = _table(L, "vip_integral_u8c3", b, 200, 97, 3, -1)
    # the no-alloc calls back to back on one stream, no host synchronisation in between, then the first frame again
    # over whatever the second call left behind
    got_a = _table(L, "vip_integral_u8c3_noalloc", a, 1337, 301, 7, -2)
    got_b = _table(L, "vip_integral_u8c3_noalloc", b, 200, 97, 3, -2)
    got_a2 = _table(L, "vip_integral_u8c3_noalloc", a, 1337, 301, 7, -2)
    torch.cuda.synchronize()
    assert torch.equal(got_a, want_a)
    assert torch.equal(got_b, want_b)
    assert torch.equal(got_a2, want_a)


def test_noalloc_4k_frame_fits(L):
    w, h, r = 3840, 2160, 15
    src = _image(w, h, seed=13)
    want, got = _tables(L, src, w, h, r)
    assert np.array_equal(got, want)


def test_noalloc_refuses_what_does_not_fit(L):
    w, h, r = 7680, 4320, 15          # needs ~63 MB of aggregates: beyond the 20 MiB array
    src = torch.zeros((16,), dtype=torch.uint8, device="cuda")      # never dereferenced: the size check comes first
    out = torch.full((16,), 5, dtype=torch.int32, device="cuda")
    assert L.vip_integral_u8c3_noalloc(src.data_ptr(), out.data_ptr(), w, h, r, None) == -2    # VIP_E_SIZE
    assert L.vip_integral_u8c3_noalloc(src.data_ptr(), out.data_ptr(), 2**30, 2**30, 2**29, None) == -2
    assert L.vip_integral_u8c3_noalloc(None, out.data_ptr(), 8, 8, 1, None) == -1              # VIP_E_NULL
    torch.cuda.synchronize()
    assert int(out.sum()) == 5 * 16
