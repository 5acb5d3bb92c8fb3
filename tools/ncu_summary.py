#!/usr/bin/env python
"""Summarise ncu output brought back in gpurun_out/ into small text files under profiles/ (tracked).

    python tools/ncu_summary.py launches gpurun_out/x_launches.csv            > profiles/rNN_x_launches.txt
    python tools/ncu_summary.py full gpurun_out/x.ncu-rep [kernel-regex]      > profiles/rNN_x_full.txt

`launches` = the per-launch `gpu__time_duration.sum` list (cold-cache, serialised: compare shares, not absolutes).
`full`     = one `ncu --set full` capture: duration, DRAM bytes, occupancy, issue / stall breakdown per launch and the
             SASS instructions with the most stall samples.
"""
import collections
import csv
import io
import re
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
    "launch__shared_mem_per_block_allocated", "launch__waves_per_multiprocessor", "launch__grid_size",
    "launch__block_size", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_active.avg",
    "smsp__average_warp_latency_per_inst_issued.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]


NOTE = None   # --note "...": what was captured (command, worlds, STEPS PER LAUNCH), printed in every summary header


def header_note():
    if NOTE:
        print(f"# captured: {NOTE}")


def launches(path):
    rows = [r for r in csv.reader(open(path, newline="")) if len(r) > 10]
    hdr = rows[0]
    ki, vi, gi, bi = (hdr.index(k) for k in ("Kernel Name", "Metric Value", "Grid Size", "Block Size"))
    agg = collections.OrderedDict()
    for r in rows[1:]:
        agg.setdefault((r[ki], r[gi], r[bi]), []).append(float(r[vi].replace(",", "")))
    total = sum(sum(v) for v in agg.values())
    print(f"# ncu --metrics gpu__time_duration.sum --clock-control none : {path}")
    header_note()
    print(f"# {sum(len(v) for v in agg.values())} launches, {total / 1e3:.1f} us total (cold-cache, serialised)")
    print(f"{'share':>7} {'n':>5} {'avg_us':>9} {'min_us':>9} {'max_us':>9}  grid block kernel")
    for (k, g, b), v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"{100 * sum(v) / total:6.1f}% {len(v):5d} {sum(v) / len(v) / 1e3:9.2f} {min(v) / 1e3:9.2f} "
              f"{max(v) / 1e3:9.2f}  {g} {b} {k[:110]}")


def ncu(*args):
    return subprocess.run(["ncu", *args], capture_output=True, text=True).stdout


def full(path, pattern=None):
    raw = list(csv.reader(io.StringIO(ncu("-i", path, "--page", "raw", "--csv"))))
    hdr, units, body = raw[0], raw[1], raw[2:]
    ki = hdr.index("Kernel Name")
    print(f"# ncu --set full --clock-control none : {path}")
    header_note()
    for r in body:
        if pattern and not re.search(pattern, r[ki]):
            continue
        print(f"\n== launch id {r[0]}: {r[ki][:120]}")
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"  {k:85s} {r[i]:>16s} {units[i]}")
    sass = list(csv.reader(io.StringIO(ncu("-i", path, "--page", "source", "--csv", "--print-source", "sass"))))
    starts = [i for i, r in enumerate(sass) if r and r[0] == "Address"]
    if not starts:
        return
    h = sass[starts[0]]
    end = starts[1] - 1 if len(starts) > 1 else len(sass)
    rows = [r for r in sass[starts[0] + 1:end] if len(r) == len(h)]
    si, ie = h.index("# Samples"), h.index("Instructions Executed")
    stall_cols = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
    tot = sum(int(r[si] or 0) for r in rows)
    print(f"\n== SASS of the first captured launch: {len(rows)} instructions, {tot} stall samples; top 25 by samples")
    print(f"{'idx':>5} {'samples':>7} {'share':>6} {'executed':>9}  top-stall          instruction")
    for i in sorted(sorted(range(len(rows)), key=lambda i: -int(rows[i][si] or 0))[:25]):
        r = rows[i]
        best = max(stall_cols, key=lambda c: int(r[c] or 0))
        print(f"{i:5d} {r[si]:>7} {100 * int(r[si] or 0) / max(tot, 1):5.1f}% {r[ie]:>9}  {h[best]:18s} {r[1].strip()[:90]}")
    by = collections.Counter()
    for r in rows:
        for c in stall_cols:
            by[h[c]] += int(r[c] or 0)
    print("\n== stall samples by reason: " + ", ".join(f"{k}={v}" for k, v in by.most_common(8)))




def lines(rep, cubin, func_pattern):
    """Join the per-SASS-instruction counters of the first captured launch with nvdisasm -g line info (by instruction
    order) and print executed warp-instructions / stall samples per source line."""
    dis = subprocess.run(["nvdisasm", "-c", "-g", cubin], capture_output=True, text=True).stdout.splitlines()
    cur, infn, where = None, False, []
    for ln in dis:
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m:
            infn = re.search(func_pattern, m.group(1)) is not None and not where
            continue
        if ln.startswith("//-----") and where:
            infn = False
        if not infn:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
        elif re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln):
            where.append(cur)
    sass = list(csv.reader(io.StringIO(ncu("-i", rep, "--page", "source", "--csv", "--print-source", "sass"))))
    starts = [i for i, r in enumerate(sass) if r and r[0] == "Address"]
    h = sass[starts[0]]
    end = starts[1] - 1 if len(starts) > 1 else len(sass)
    rows = [r for r in sass[starts[0] + 1:end] if len(r) == len(h)]
    si, ie = h.index("# Samples"), h.index("Instructions Executed")
    if len(rows) != len(where):
        print(f"# warning: {len(rows)} profiled instructions vs {len(where)} disassembled")
    agg = collections.OrderedDict()
    for r, w in zip(rows, where):
        a = agg.setdefault(w, [0, 0])
        a[0] += int(r[ie] or 0)
        a[1] += int(r[si] or 0)
    launches_ = max(int(r[ie] or 0) for r in rows[:3]) or 1
    tot_i, tot_s = sum(a[0] for a in agg.values()), sum(a[1] for a in agg.values())
    print(f"# per source line: warp-instructions per warp (executed / {launches_} warps), stall samples; totals "
          f"{tot_i / launches_:.0f} instr/warp, {tot_s} samples")
    for (f, l), (n, s) in sorted(agg.items(), key=lambda kv: (kv[0][0], kv[0][1])):
        if n / launches_ >= 4 or s >= 4:
            print(f"{f:>16}:{l:<4} {n / launches_:8.1f} instr {s:5d} samples")


if __name__ == "__main__":
    if "--note" in sys.argv:
        i = sys.argv.index("--note")
        NOTE = sys.argv[i + 1]
        del sys.argv[i:i + 2]
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    elif sys.argv[1] == "lines":
        lines(sys.argv[2], sys.argv[3], sys.argv[4])
    else:
        full(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)
