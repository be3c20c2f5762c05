"""Summaries of ncu output for profiles/: (1) launch list CSV -> per-kernel totals and shares;
(2) raw-page CSV of a --set full capture -> one line per kernel launch with the metrics the roofline
uses (duration, DRAM bytes, issue-active, occupancy, L1/L2 hit rates, registers).
usage: python tools/ncu_summary.py launches LAUNCHES.csv | raw FRAME_RAW.csv"""
import csv
import sys
from collections import defaultdict


def launches(path):
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr = rows[0]
    iN, iV, iM = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name")
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows[1:]:
        if r[iM] != "gpu__time_duration.sum":
            continue
        v = float(r[iV].replace(",", ""))
        unit = r[hdr.index("Metric Unit")]
        v = v / 1000.0 if unit in ("ns", "nsecond") else v * (1000.0 if unit in ("ms", "msecond") else 1.0)
        tot[r[iN]] += v
        cnt[r[iN]] += 1
    T = sum(tot.values())
    print("# us_total launches us_each share kernel   (total %.1f us; cold-cache serialised times: compare SHARES)" % T)
    for k in sorted(tot, key=lambda k: -tot[k]):
        print("%10.1f %4d %8.1f %5.1f%% %s" % (tot[k], cnt[k], tot[k] / cnt[k], 100 * tot[k] / T, k[:150]))


WANT = [
    ("gpu__time_duration.sum", "dur"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("sm__maximum_warps_per_active_cycle_pct", "occ_theo%"),
    ("l1tex__t_sector_hit_rate.pct", "l1hit%"),
    ("lts__t_sector_hit_rate.pct", "l2hit%"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem_wavefronts"),
    ("smsp__inst_executed.sum", "warp_inst"),
    ("sm__inst_executed_pipe_fma.sum", "pipe_fma"),
    ("sm__inst_executed_pipe_alu.sum", "pipe_alu"),
    ("sm__inst_executed_pipe_xu.sum", "pipe_xu"),
    ("sm__inst_executed_pipe_lsu.sum", "pipe_lsu"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu%"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"),
]


def raw(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    iK = hdr.index("Kernel Name")
    cols = [(hdr.index(m), s) for m, s in WANT if m in hdr]
    print("# one line per launch of the --set full frame; bytes in MB, dur in us")
    print("# " + " ".join(s for _, s in cols) + " kernel")
    for r in rows[2:]:
        out = []
        for i, s in cols:
            v = r[i].replace(",", "")
            try:
                f = float(v)
            except ValueError:
                out.append(v)
                continue
            u = units[i]
            if s == "dur":
                f = f / 1000.0 if u in ("ns", "nsecond") else f * (1000.0 if u in ("ms", "msecond") else 1.0)
            if s.startswith("dram_"):
                f = f * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1e-6)
            out.append("%.1f" % f if abs(f) < 1e7 else "%.3g" % f)
        print(" ".join("%9s" % o for o in out) + "  " + r[iK][:60])


if __name__ == "__main__":
    {"launches": launches, "raw": raw}[sys.argv[1]](sys.argv[2])
