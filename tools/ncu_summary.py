#!/usr/bin/env python
"""Development aid: one-page text summary of an `ncu --set full` capture from its exported pages.

    python tools/ncu_summary.py <dir> <name> [algorithmic bytes per launch]   (reads <dir>/<name>.raw.csv, .details.txt, .source_sass.csv.gz)

Prints duration, DRAM bytes (read + write) against the algorithmic bytes, registers / occupancy, issue-slot and pipe
utilisation, the warp-stall breakdown per issued instruction, and the opcode mix of the hottest instructions of the source
page.  The summaries under profiles/ are this tool's output."""
import collections
import csv
import gzip
import os
import sys


def num(s):
    try:
        return float(s.replace(",", ""))
    except (ValueError, AttributeError):
        return None


def main():
    d, name = sys.argv[1], sys.argv[2]
    algo = float(sys.argv[3]) if len(sys.argv) > 3 else None
    rows = list(csv.reader(open(os.path.join(d, name + ".raw.csv"))))
    hdr, units, vals = rows[0], rows[1], rows[2]
    m = dict(zip(hdr, vals))
    u = dict(zip(hdr, units))
    print(f"# ncu --set full --clock-control none: {m.get('Kernel Name', name)}")
    print(f"# grid {m.get('Grid Size')} block {m.get('Block Size')}   source: tools/ncu_summary.py {name}")

    def show(key, label=None, scale=1.0, fmt="{:.4g}"):
        v = num(m.get(key))
        if v is not None:
            print(f"{label or key:58s} {fmt.format(v * scale)} {u.get(key, '')}")
        return v

    dur = show("gpu__time_duration.sum", "duration")
    rd, wr = num(m.get("dram__bytes_read.sum")), num(m.get("dram__bytes_write.sum"))
    if rd is not None and wr is not None:
        mult = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}
        rdb, wrb = rd * mult.get(u["dram__bytes_read.sum"], 1.0), wr * mult.get(u["dram__bytes_write.sum"], 1.0)
        print(f"{'dram bytes read + written':58s} {rdb:.6g} + {wrb:.6g} = {rdb + wrb:.6g} B")
        if algo:
            print(f"{'  / algorithmic bytes of the launch':58s} {(rdb + wrb) / algo:.4f}  (algorithmic {algo:.6g} B)")
    for key, label in (("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput, % of peak"),
                       ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
                       ("launch__registers_per_thread", "registers / thread"),
                       ("launch__shared_mem_per_block_dynamic", "dynamic shared memory / CTA"),
                       ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
                       ("sm__maximum_warps_per_active_cycle_pct", "theoretical occupancy %"),
                       ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
                       ("smsp__inst_executed.sum", "warp instructions executed"),
                       ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "pipe fma (FFMA/FMUL/FADD issue) %"),
                       ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "pipe fma cycles active %"),
                       ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "pipe alu %"),
                       ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "pipe xu (conversions, transcendentals) %"),
                       ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "pipe lsu %"),
                       ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "shared-memory wavefronts % of peak"),
                       ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "shared-memory bank conflicts")):
        show(key, label)
    print("warp stalls per issued instruction (smsp__average_warps_issue_stalled_*_per_issue_active):")
    st = []
    for k in hdr:
        if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio") and "not_issued" not in k:
            v = num(m[k])
            if v and v > 0.03:
                st.append((v, k[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
    for v, k in sorted(st, reverse=True):
        print(f"    {k:28s} {v:.3f}")
    src = os.path.join(d, name + ".source_sass.csv.gz")
    if os.path.exists(src):
        rows = list(csv.reader(gzip.open(src, "rt")))
        h2, body = rows[1], rows[2:]
        ix = {c: i for i, c in enumerate(h2)}
        tot = sum(int(r[ix["# Samples"]] or 0) for r in body)
        execd = collections.Counter()
        ops = collections.defaultdict(lambda: [0, 0])
        for r in body:
            t = r[ix["Source"]].split()
            op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
            n = int(r[ix["Instructions Executed"]] or 0)
            ops[op][0] += n
            ops[op][1] += int(r[ix["# Samples"]] or 0)
            execd[op] += n
        allx = sum(execd.values())
        print(f"source page: {len(body)} SASS instructions, {tot} samples; executed warp instructions by opcode (share, share of samples):")
        for op, (n, s) in sorted(ops.items(), key=lambda kv: -kv[1][0])[:12]:
            print(f"    {op:10s} {100.0 * n / allx:5.1f} %   {100.0 * s / max(tot, 1):5.1f} %")


if __name__ == "__main__":
    main()
