#!/usr/bin/env python
"""Turn ncu output (read on the CPU box) into the small text/JSON summaries committed under profiles/.

  python tools/ncu_summary.py launches <launches.csv> [...]      per-kernel launch counts, mean time, share of the step
  python tools/ncu_summary.py full <prof.ncu-rep> [...]          key metrics of each profiled launch (ncu --set full)
  python tools/ncu_summary.py traffic <name>=<prof.ncu-rep> ...  JSON {name: dram bytes per launch} for bench.py
  python tools/ncu_summary.py constants <key>=<prof.ncu-rep>:<pixels>:<taps per pixel> ...
                                                                 profiles/ncu_constants.json: what bench.py reports as
                                                                 issue.util_ncu / roofline.traffic (hardware counters of the
                                                                 FINAL tree's kernels, never literals in bench.py)
"""
import collections
import csv
import json
import subprocess
import sys

FULL_METRICS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs/thread"),
    ("launch__shared_mem_per_block_dynamic", "dyn smem/block"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "pipe FMA %"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "pipe ALU %"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "pipe XU %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "pipe LSU %"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "pipe FP64 %"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "shared wavefronts"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "shared bank conflicts"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram throughput %"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram throughput %"),
    ("lts__t_bytes.sum", "L2 bytes"),
]
STALL_PREFIX = "smsp__average_warps_issue_stalled_"


def raw_rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(out.splitlines()))
    return rows[0], rows[1], rows[2:]


def to_bytes(value, unit):
    v = float(value.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def cmd_launches(paths):
    for path in paths:
        rows = [r for r in csv.reader(open(path)) if len(r) > 10]
        hdr = rows[0]
        ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
        agg = collections.OrderedDict()
        for r in rows[1:]:
            v = float(r[vi].replace(",", ""))
            v = {"ns": v / 1e3, "us": v, "ms": v * 1e3, "s": v * 1e6}.get(r[ui], v)
            agg.setdefault(r[ki], []).append(v)
        tot = sum(sum(v) for v in agg.values())
        print(f"# {path}: {sum(len(v) for v in agg.values())} launches, {tot:.1f} us total (ncu: cold cache, serialised -> compare shares)")
        for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
            print(f"{sum(v) / tot:6.3f}  n={len(v):4d}  mean {sum(v) / len(v):10.1f} us  {k[:110]}")


def cmd_full(paths):
    for path in paths:
        hdr, units, rows = raw_rows(path)
        for v in rows:
            print(f"# {path}: {v[hdr.index('Kernel Name')][:100]}")
            for name, label in FULL_METRICS:
                if name in hdr:
                    i = hdr.index(name)
                    print(f"  {label:26s} {v[i]:>18s} {units[i]:14s} ({name})")
            stalls = []
            for i, n in enumerate(hdr):
                if n.startswith(STALL_PREFIX) and n.endswith("_per_issue_active.ratio") and "not_issued" not in n:
                    try:
                        stalls.append((float(v[i]), n[len(STALL_PREFIX):-len("_per_issue_active.ratio")]))
                    except ValueError:
                        pass
            stalls.sort(reverse=True)
            print("  stall reasons (warps per issue-active cycle): " + ", ".join(f"{n} {x:.2f}" for x, n in stalls[:8]))


def cmd_traffic(specs):
    out = {}
    for spec in specs:
        name, path = spec.split("=", 1)
        hdr, units, rows = raw_rows(path)
        tot = 0.0
        for v in rows:
            for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                i = hdr.index(m)
                tot += to_bytes(v[i], units[i])
        out[name] = tot / max(1, len(rows))
    print(json.dumps(out, indent=1))


CONST_METRICS = {
    "gpu__time_duration.sum": "duration",
    "smsp__inst_executed.sum": "warp_inst_executed",
    "smsp__inst_issued.sum": "warp_inst_issued",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active": "pipe_fma_cycles_active_pct",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active": "pipe_fmaheavy_cycles_active_pct",
    "sm__pipe_fmalite_cycles_active.avg.pct_of_peak_sustained_active": "pipe_fmalite_cycles_active_pct",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active": "inst_pipe_fma_pct",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "inst_pipe_alu_pct",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active": "inst_pipe_xu_pct",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active": "inst_pipe_lsu_pct",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum": "shared_wavefronts",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": "shared_bank_conflicts",
    "launch__registers_per_thread": "registers_per_thread",
}


def cmd_constants(specs):
    out = {}
    for spec in specs:
        key, rest = spec.split("=", 1)
        path, pixels, taps = rest.rsplit(":", 2)
        hdr, units, rows = raw_rows(path)
        v = rows[0]
        d = {"source": path.split("/")[-1], "kernel": v[hdr.index("Kernel Name")][:80], "launches_in_report": len(rows)}
        for m, label in CONST_METRICS.items():
            if m in hdr:
                i = hdr.index(m)
                d[label] = float(v[i].replace(",", ""))
                if label == "duration":
                    d["duration_unit"] = units[i]
        d["dram_bytes_per_launch"] = sum(to_bytes(v[hdr.index(m)], units[hdr.index(m)])
                                         for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        n_taps = float(pixels) * float(taps)
        d["pixels"], d["taps_per_pixel"] = int(pixels), int(taps)
        d["thread_instr_per_tap"] = d["warp_inst_executed"] * 32 / n_taps      # whole kernel: staging, prologue, epilogue included
        stalls = []
        for i, n in enumerate(hdr):
            if n.startswith(STALL_PREFIX) and n.endswith("_per_issue_active.ratio") and "not_issued" not in n:
                try:
                    stalls.append((float(v[i]), n[len(STALL_PREFIX):-len("_per_issue_active.ratio")]))
                except ValueError:
                    pass
        d["top_stalls_per_issue_active"] = {n: round(x, 2) for x, n in sorted(stalls, reverse=True)[:5]}
        out[key] = d
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    {"launches": cmd_launches, "full": cmd_full, "traffic": cmd_traffic, "constants": cmd_constants}[sys.argv[1]](sys.argv[2:])
