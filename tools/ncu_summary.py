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



STAGE_OF = [  # kernel-name fragment -> bench.py stage
    ("lbs_forward_kernel", "lbs_forward"), ("lbs_backward_kernel", "lbs_backward"),
    ("preprocess_backward_kernel", "preprocess_backward"), ("preprocess_kernel", "preprocess"),
    ("radix_", "sort_pairs"), ("inclusive_scan_kernel", "inclusive_scan"), ("duplicate_", "duplicate_with_keys"),
    ("tile_ranges", "tile_split"), ("l2_", "tile_split"),
    ("render_forward_kernel", "render_forward"), ("render_backward_kernel", "render_backward"),
]


def traffic(path, source):
    """profiles/ncu_dram_traffic.json (read by bench.py for roofline.traffic / issue_frac) from the raw page of
    the one-frame --set full capture: per stage, DRAM bytes and duration summed over its launches, issue-active
    duration-weighted."""
    import json
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    col = {m: hdr.index(m) for m in ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum",
                                     "dram__bytes_write.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active")}

    def num(r, m, scale):
        i = col[m]
        return float(r[i].replace(",", "")) * scale.get(units[i], 1.0)

    by, us, iss = defaultdict(float), defaultdict(float), defaultdict(float)
    for r in rows[2:]:
        stage = next((s for frag, s in STAGE_OF if frag in r[col["Kernel Name"]]), None)
        if stage is None:
            continue
        t = num(r, "gpu__time_duration.sum", {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3})
        b = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        by[stage] += num(r, "dram__bytes_read.sum", b) + num(r, "dram__bytes_write.sum", b)
        us[stage] += t
        iss[stage] += t * num(r, "smsp__issue_active.avg.pct_of_peak_sustained_active", {})
    out = {"source": source,
           "bytes_per_frame": {s: int(round(v, -5)) for s, v in by.items()},
           "issue_active_pct": {s: round(iss[s] / us[s], 1) for s in us},
           "issue_source": "same capture, smsp__issue_active.avg.pct_of_peak_sustained_active (duration-weighted over "
                           "the stage's launches)",
           "ncu_us_per_frame": {s: round(v, 1) for s, v in us.items()}}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    if sys.argv[1] == "traffic":
        traffic(sys.argv[2], sys.argv[3])
    else:
        {"launches": launches, "raw": raw}[sys.argv[1]](sys.argv[2])
