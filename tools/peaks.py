#!/usr/bin/env python
"""Measured issue-rate roofs of the bilateral stencils on this GPU (tools/probe/issue_probe.cu through ctypes).
  python tools/peaks.py [out.json]
Prints one JSON object: per instruction kind the measured rate in thread-instructions/s, the same per SM per clock at the
SM clock sampled under load, and the nominal 128 / clk / SM figure next to it."""
import ctypes as C
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "tools", "probe", "libvip_probe.so")
KINDS = {0: "ffma", 1: "ffma2_packed", 2: "ffma_plus_vabsdiff4", 3: "lds32", 4: "i2f_u8", 5: "ffma2_plus_vabsdiff4",
         6: "ffma2_plus_lds32", 7: "ffma_plus_lop3", 8: "ffma2_plus_ffma"}


def load():
    if not os.path.exists(LIB):
        return None
    lib = C.CDLL(LIB)
    lib.vip_probe_issue_rate.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_float)]
    lib.vip_probe_issue_rate.restype = C.c_int
    return lib


def measure(lib, kind, trips=4000, repeats=3):
    """best of `repeats` (the first run also warms the clocks up); returns (thread-instr/s, ms)"""
    best = (0.0, 0.0)
    for _ in range(repeats):
        rate, ms = C.c_double(0), C.c_float(0)
        rc = lib.vip_probe_issue_rate(kind, trips, C.byref(rate), C.byref(ms))
        if rc != 0:
            raise RuntimeError(f"vip_probe_issue_rate({kind}) failed: cudaError {rc}")
        if rate.value > best[0]:
            best = (rate.value, ms.value)
    return best


def sm_clock_mhz():
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm", "--format=csv,noheader,nounits", "-i", "0"],
                             capture_output=True, text=True, timeout=10).stdout.strip().split(",")
        return float(out[0]), float(out[1])
    except Exception:
        return None, None


def main():
    lib = load()
    if lib is None:
        sys.exit("tools/probe/libvip_probe.so is missing: run python -c 'import __graft_entry__ as g; g.build()'")
    import torch
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    res = {"device": torch.cuda.get_device_name(0), "sm_count": sms, "kinds": {}}
    for kind, name in KINDS.items():
        rate, ms = measure(lib, kind)
        cur, mx = sm_clock_mhz()
        clk = (mx or 1965.0) * 1e6
        res["kinds"][name] = {"thread_instr_per_s": rate, "ms": ms, "per_sm_per_clk_at_max_clock": rate / sms / clk,
                              "sm_mhz_after": cur}
    res["sm_max_mhz"] = sm_clock_mhz()[1]
    res["nominal_thread_instr_per_s"] = sms * 128 * (res["sm_max_mhz"] or 1965.0) * 1e6
    line = json.dumps(res)
    print(line)
    if len(sys.argv) > 1:
        with open(sys.argv[1], "w") as f:
            f.write(json.dumps(res, indent=1) + "\n")


if __name__ == "__main__":
    main()
