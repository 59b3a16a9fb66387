#!/usr/bin/env python
"""Summarise an .ncu-rep: key metrics per kernel + top stalled SASS lines.
  python tools/ncu_summary.py gpurun_out/prof.ncu-rep [kernel-regex] [ntop]"""
import csv
import io
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'smsp__inst_executed_op_shared_atom.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum', 'lts__t_sectors_op_atom.sum',
        'lts__t_sectors_op_red.sum', 'sm__cycles_elapsed.max']


def run(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    rx = sys.argv[2] if len(sys.argv) > 2 else None
    ntop = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    rows = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        name = r[hdr.index('Kernel Name')]
        print("== kernel:", name[:100])
        for k in KEYS:
            if k in hdr:
                print("  %-70s %s %s" % (k, r[hdr.index(k)], units[hdr.index(k)]))
        st = []
        for i, h in enumerate(hdr):
            if 'issue_stalled' in h and h.endswith('per_issue_active.ratio') is False and \
                    'pcsamp_warps_issue_stalled' in h and 'not_issued' not in h:
                try:
                    st.append((float(r[i]), h.replace('smsp__pcsamp_warps_issue_stalled_', '')))
                except ValueError:
                    pass
        tot = sum(v for v, _ in st) or 1
        print("  stalls:", ", ".join("%s %.0f%%" % (h, 100 * v / tot) for v, h in sorted(st, reverse=True)[:7]))
    args = ["-i", rep, "--page", "source", "--csv", "--print-source", "sass"]
    if rx:
        args += ["--kernel-name", "regex:" + rx]
    rows = list(csv.reader(io.StringIO(run(args))))
    if len(rows) < 3:
        return
    hdr = rows[1]
    iS, iN, iE = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    cols = [c for c in ("stall_long_sb", "stall_short_sb", "stall_mio", "stall_wait", "stall_math",
                        "stall_barrier", "stall_lg", "stall_branch_resolving", "stall_not_selected")
            if c in hdr]
    data = []
    for r in rows[2:]:
        if len(r) < len(hdr) or r[0] == "Address":
            break
        data.append(r)
    tot = sum(int(r[iN]) for r in data) or 1
    print("== top stalled SASS of the first matching kernel (%d instructions, %d samples)" % (len(data), tot))
    top = sorted(range(len(data)), key=lambda i: -int(data[i][iN]))[:ntop]
    for i in sorted(top):
        r = data[i]
        why = " ".join("%s=%s" % (c.replace("stall_", ""), r[hdr.index(c)]) for c in cols if r[hdr.index(c)] not in ("0", ""))
        print("  %5d %-64s %5.1f%% exec=%s %s" % (i, r[iS].strip()[:64], 100.0 * int(r[iN]) / tot, r[iE], why))


if __name__ == "__main__":
    main()
