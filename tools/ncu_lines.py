"""Per-source-line instruction / stall-sample totals of one kernel from an .ncu-rep.
usage: python tools/ncu_lines.py REPORT.ncu-rep KERNEL_REGEX CUBIN [min_pct] [MANGLED_REGEX] [NAME_SUBSTR]
(KERNEL_REGEX matches ncu's demangled name, MANGLED_REGEX the symbol in the cubin; first launch only)
Joins ncu's SASS page (per-instruction counts) with nvdisasm -g line info by instruction offset."""
import csv, io, re, subprocess, sys
from collections import defaultdict

rep, kre, cubin = sys.argv[1:4]
min_pct = float(sys.argv[4]) if len(sys.argv) > 4 else 0.3
mre = sys.argv[5] if len(sys.argv) > 5 else kre
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# one table per launch: "Kernel Name" row, header row, instruction rows; take the first launch whose
# demangled name contains NAME_SUBSTR (argv[6], optional)
sub = sys.argv[6] if len(sys.argv) > 6 else ""
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
pick = next(i for i in starts if sub in rows[i][1])
hdr = rows[pick + 1]
iA, iS, iI, iN = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
iT = hdr.index("Predicated-On Thread Instructions Executed")
body = []
for r in rows[pick + 2:]:
    if r and r[0] == "Kernel Name":
        break
    if len(r) > iN and r[iA].startswith("0x"):
        body.append(r)
base = int(body[0][iA], 16)
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
line_of, cur, infn = {}, None, False
for l in dis:
    if l.startswith(".text.") and re.search(mre, l): infn = True; continue
    if l.startswith(".text.") and infn and not re.search(mre, l): infn = False
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", l)
    if m and infn: line_of[int(m.group(1), 16)] = cur
inst, samp, thr = defaultdict(int), defaultdict(int), defaultdict(int)
for r in body:
    key = line_of.get(int(r[iA], 16) - base, ("?", 0))
    inst[key] += int(r[iI]); samp[key] += int(r[iN]); thr[key] += int(r[iT])
ti, ts = sum(inst.values()), sum(samp.values())
print("total warp instructions %d, samples %d" % (ti, ts))
for k in sorted(inst, key=lambda k: (k[0], k[1])):
    if 100 * inst[k] / ti >= min_pct or 100 * samp[k] / max(ts, 1) >= min_pct:
        print("%-24s %5d  inst %5.2f%%  samples %5.2f%%  lanes %4.1f" % (k[0], k[1], 100 * inst[k] / ti, 100 * samp[k] / max(ts, 1), thr[k] / max(inst[k], 1)))
