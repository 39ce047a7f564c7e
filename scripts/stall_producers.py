"""Attribute long-scoreboard stall samples of one kernel to the loads that caused them.

usage: python scripts/stall_producers.py <report.ncu-rep> <object.o> <mangled-kernel-substring>

The ncu source page (SASS view) reports a stall at the instruction that waits; this walks the
instruction stream, remembers which load last wrote each register, and charges the waiting
samples to that load's kind (LDL/LDG/LDS/...) and source line (nvdisasm line info of the same
function in the object file; -lineinfo build).  Straight-line approximation, good enough to
rank lines.
"""
import collections, csv, os, re, subprocess, sys, tempfile

rep, obj, pat = sys.argv[1:4]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, check=True, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "--print-line-info", cubin], cwd=tmp, capture_output=True, text=True).stdout
lines, cur, infn = [], None, False
for line in sass.splitlines():
    if line.startswith(".text."):
        infn = pat in line
        continue
    if not infn:
        continue
    if line.lstrip().startswith(".section"):
        infn = False
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m:
        lines.append((cur, m.group(2).strip()))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
print(rows[0][1][:100])
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
if len(lines) != len(data):
    print(f"warning: {len(lines)} instructions in the object, {len(data)} in the report")
tot = collections.Counter()
for r in data:
    for k in hdr:
        if k.startswith("stall_") and "Not Issued" not in k:
            tot[k] += int(r[ix[k]] or 0)
    tot["samples"] += int(r[ix["# Samples"]] or 0)
print({k: v for k, v in tot.most_common(9)})
lastw, prod, cons = {}, collections.Counter(), collections.Counter()
for (ln, _), r in zip(lines, data):
    s = r[ix["Source"]].strip()
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)\s*(.*)", s)
    if not m:
        continue
    op, args = m.group(2), m.group(3)
    lsb = int(r[ix["stall_long_sb"]] or 0)
    regs = re.findall(r"\bR(\d+)\b", args)
    if lsb:
        srcs = regs[1:] if not op.startswith(("ST", "RED", "ATOM")) else regs
        ps = set(lastw[x] for x in srcs if lastw.get(x))
        if ps:
            for p in ps:
                prod[p] += lsb / len(ps)
        else:
            cons[("?", ln, op)] += lsb
    if regs and not op.startswith(("ST", "RED", "BRA", "EXIT")):
        d, base = regs[0], op.split(".")[0]
        if base in ("LDL", "LDG", "ATOMG", "LDS", "LDC", "LD"):
            width = 4 if ".128" in op else 2 if ".64" in op else 1
            for k in range(width):
                lastw[str(int(d) + k)] = (base, ln)
        else:
            lastw[d] = None
byf = collections.Counter()
for (b, ln), v in prod.items():
    byf[(b, ln[0] if ln else None)] += v
print("-- by load kind and file")
for k, v in byf.most_common(12):
    print(k, round(v))
print("-- by producing line")
for k, v in prod.most_common(30):
    print(k, round(v))
print("-- unattributed (waiting instruction)")
for k, v in cons.most_common(10):
    print(k, v)
# instruction and sample share per file / hottest lines (same attribution)
inst, samp, thr = collections.Counter(), collections.Counter(), collections.Counter()
for (ln, _), r in zip(lines, data):
    n = int(r[ix["Instructions Executed"]] or 0)
    inst[ln] += n
    samp[ln] += int(r[ix["# Samples"]] or 0)
    thr[ln] += int(r[ix["Thread Instructions Executed"]] or 0)
ti, ts = sum(inst.values()), sum(samp.values())
fi, fs = collections.Counter(), collections.Counter()
for ln in inst:
    fi[ln[0] if ln else None] += inst[ln]
    fs[ln[0] if ln else None] += samp[ln]
print("-- per file: inst%  samples%")
for k, v in fi.most_common(14):
    print(f"{str(k):28s} {100 * v / ti:5.1f} {100 * fs[k] / ts:5.1f}")
print("-- hottest lines by samples: samples% inst% avg threads")
for ln, v in samp.most_common(40):
    print(f"{str(ln):34s} {100 * v / ts:5.1f} {100 * inst[ln] / ti:5.1f} {thr[ln] / max(inst[ln], 1):5.1f}")
