#!/usr/bin/env python
"""Turns ncu output brought back in gpurun_out/ into the small text summaries committed under profiles/.

    python profiles/summarize_ncu.py launches gpurun_out/launches.csv              > profiles/rNN_launches.txt
    python profiles/summarize_ncu.py kernel   gpurun_out/prof.ncu-rep [kernel-idx] > profiles/rNN_<kernel>.txt

    python profiles/summarize_ncu.py json     gpurun_out/prof.ncu-rep configN <segments> <committed summary path> [idx]
                                              -> updates profiles/ncu_summary.json (what bench.py quotes as `executed`,
                                                 `traffic`, `issue_active`, `lanes_per_inst` for that configuration)

`launches`: per-kernel totals and SHARES of a `--metrics gpu__time_duration.sum` pass (cold-cache, serialised: the
shares are what counts).  `kernel`: the counters the north star names (warp execution efficiency, L1/L2 hit rates,
issue rate, DRAM bytes and GB/s), registers/occupancy, the stall-reason mix, the opcode mix and the per-instruction
active-thread histogram from the source page.
"""
from __future__ import annotations

import collections
import csv
import io
import subprocess
import sys

RAW_KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__thread_inst_executed_per_inst_executed.pct", "sm__inst_executed.avg.per_cycle_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_alu.sum.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.sum.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.sum.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.sum.pct_of_peak_sustained_active",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
    "dram__bytes_write.sum.per_second", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed.sum", "sm__cycles_elapsed.avg.per_second",
    "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum.per_cycle_elapsed",
    "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum.per_cycle_elapsed",
    "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum.per_cycle_elapsed",
    "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum",
    "smsp__average_warp_latency_per_inst_issued.ratio",
]


def ncu_csv(report, page):
    out = subprocess.run(["ncu", "-i", report, "--page", page, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    total, count = collections.defaultdict(float), collections.Counter()
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        ms = v * {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0,
                  "s": 1e3, "second": 1e3}.get(unit, 1.0)
        name = r["Kernel Name"]
        total[name] += ms
        count[name] += 1
    grand = sum(total.values())
    print(f"# {path}: {sum(count.values())} launches, {grand:.3f} ms of device time (cold-cache, serialised)")
    print(f"{'kernel':80s} {'launches':>8s} {'total ms':>12s} {'avg ms':>10s} {'share':>8s}")
    for name, ms in sorted(total.items(), key=lambda kv: -kv[1]):
        print(f"{name[:80]:80s} {count[name]:8d} {ms:12.3f} {ms / count[name]:10.3f} {100 * ms / grand:7.2f}%")


def kernel(report, index=0):
    rows = ncu_csv(report, "raw")
    header, units = rows[0], rows[1]
    values = rows[2 + index]
    col = {h: i for i, h in enumerate(header)}
    print(f"# {report}: launch {index}: {values[col['Kernel Name']]}  (ncu --set full, --clock-control none)")
    for key in RAW_KEYS:
        if key in col:
            print(f"{key:75s} {values[col[key]]:>18s} {units[col[key]]}")
    print("\n# warp stall reasons (warps stalled per issued instruction; 'selected' = 1 is the issuing warp)")
    def num(v):
        try:
            return float(str(v).replace(",", ""))
        except ValueError:
            return 0.0
    stalls = [(h, values[i]) for h, i in col.items()
              if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
    for h, v in sorted(stalls, key=lambda kv: -num(kv[1]))[:8]:
        name = h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]
        print(f"  {name:28s} {num(v):6.2f}")

    src = ncu_csv(report, "source")
    # the first kernel's table: row 0 = kernel name, row 1 = header
    hdr = src[1]
    ix = {h: i for i, h in enumerate(hdr)}
    data = []
    for r in src[2:]:
        if len(r) < len(hdr) or not r[ix["Instructions Executed"]].isdigit():
            break
        data.append(r)
    tot_i = sum(int(r[ix["Instructions Executed"]]) for r in data)
    tot_t = sum(int(r[ix["Thread Instructions Executed"]]) for r in data)
    print(f"\n# source page: {len(data)} SASS instructions, {tot_i} warp instructions executed, "
          f"{tot_t / max(tot_i, 1):.2f} active threads per warp instruction")
    hist = collections.Counter()
    for r in data:
        hist[int(float(r[ix["Avg. Threads Executed"]])) // 4 * 4] += int(r[ix["Instructions Executed"]])
    for k in sorted(hist):
        print(f"  active threads {k:2d}-{k + 3:2d}: {100 * hist[k] / tot_i:5.1f}% of warp instructions")
    ops = collections.Counter()
    for r in data:
        words = r[ix["Source"]].split()
        op = words[1] if words[0].startswith("@") else words[0]
        ops[op.split(".")[0]] += int(r[ix["Instructions Executed"]])
    print("# opcode mix (% of warp instructions): " +
          ", ".join(f"{k} {100 * v / tot_i:.1f}" for k, v in ops.most_common(16)))


def to_json(report, key, segments, source, index=0):
    import json
    from pathlib import Path

    rows = ncu_csv(report, "raw")
    header, units = rows[0], rows[1]
    values = rows[2 + index]
    col = {h: i for i, h in enumerate(header)}

    def val(name, scale_units=None):
        if name not in col:
            return None
        v = float(values[col[name]].replace(",", ""))
        if scale_units:
            v *= scale_units.get(units[col[name]], 1.0)
        return v

    byte_units = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    time_units = {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}
    hz_units = {"hz": 1.0, "Khz": 1e3, "Mhz": 1e6, "Ghz": 1e9}
    ms = val("gpu__time_duration.sum", time_units)
    cycles = ms * 1e-3 * val("sm__cycles_elapsed.avg.per_second", hz_units)
    flop = (2 * val("smsp__sass_thread_inst_executed_op_ffma_pred_on.sum.per_cycle_elapsed") +
            val("smsp__sass_thread_inst_executed_op_fadd_pred_on.sum.per_cycle_elapsed") +
            val("smsp__sass_thread_inst_executed_op_fmul_pred_on.sum.per_cycle_elapsed")) * cycles
    warp_inst = val("smsp__inst_executed.sum")
    local = (val("smsp__inst_executed_op_local_ld.sum") or 0.0) + (val("smsp__inst_executed_op_local_st.sum") or 0.0)
    entry = {
        "source": source, "kernel": values[col["Kernel Name"]], "n_gpus": 1, "segments": int(segments),
        "duration_ms": ms,
        "dram_bytes": val("dram__bytes_read.sum", byte_units) + val("dram__bytes_write.sum", byte_units),
        "flop_executed": flop, "warp_inst": warp_inst,
        "lanes_per_inst": val("smsp__thread_inst_executed_per_inst_executed.ratio"),
        "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "l1_hit_pct": val("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": val("lts__t_sector_hit_rate.pct"),
        "local_ld_st_pct_of_warp_inst": 100.0 * local / warp_inst if warp_inst else None,
    }
    path = Path(__file__).resolve().parent / "ncu_summary.json"
    table = json.loads(path.read_text()) if path.exists() else {}
    table[key] = entry
    path.write_text(json.dumps(table, indent=1) + "\n")
    print(json.dumps(entry, indent=1))


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    elif sys.argv[1] == "json":
        to_json(sys.argv[2], sys.argv[3], sys.argv[4], sys.argv[5], int(sys.argv[6]) if len(sys.argv) > 6 else 0)
    else:
        kernel(sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 0)
