"""Attribute the warp-stall samples of one `ncu --set full --import-source on` capture to source files and functions.

ncu's CLI exports the source page per SASS instruction (`--page source --csv`); nvdisasm -g on the cubin inside libbcp.so
gives the file:line of every instruction of the same function.  Joined on the instruction offset, the samples are summed
per file and per enclosing function (the definition whose line range holds the line; inlined code keeps the line of its own
source).  Usage:  python scripts/ncu_by_source.py gpurun_out/x.ncu-rep k_prepIfLb1E [libbcp.so]"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def function_ranges(path):
    """[(first line, name)] of the function definitions of a source file, by a crude scan"""
    out = []
    pat = re.compile(r"^(?:template\s*<[^>]*>\s*)?(?:BCP_HD(?:_NOINLINE)?|__global__|__device__|static|inline)\b.*?\b([A-Za-z_]\w*)\s*\(")
    try:
        lines = open(path).read().splitlines()
    except OSError:
        return out
    for i, l in enumerate(lines, 1):
        m = pat.match(l.strip())
        if m and not l.strip().startswith("//"):
            out.append((i, m.group(1)))
    return out


def main():
    rep, kernel_pat = sys.argv[1], sys.argv[2]
    lib = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "cartpoleplusplus_b200", "libbcp.so")
    rows = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout.splitlines()))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    cols = {name: k for k, name in enumerate(rows[hdr])}
    inst = [r for r in rows[hdr + 1:] if len(r) > cols["# Samples"]]
    base = int(inst[0][0], 16)
    with tempfile.TemporaryDirectory() as td:
        subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=td, capture_output=True)
        dis = None
        for f in sorted(os.listdir(td)):
            txt = subprocess.run(["nvdisasm", "-g", os.path.join(td, f)], capture_output=True, text=True).stdout
            m = re.search(r"^\.text\.(\S*%s\S*):$" % re.escape(kernel_pat), txt, re.M)
            if m:
                start = m.end()
                nxt = re.search(r"^\.text\.", txt[start:], re.M)
                dis = txt[start:start + nxt.start()] if nxt else txt[start:]
                break
    if dis is None:
        raise SystemExit("kernel not found in " + lib)
    where, cur = {}, ("?", 0)
    for l in dis.splitlines():
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+\S", l)
        if m:
            where[int(m.group(1), 16)] = cur
    by_file, by_func, by_line = collections.Counter(), collections.Counter(), collections.Counter()
    inst_by_func = collections.Counter()
    ranges = {}
    total = 0
    for r in inst:
        off = int(r[0], 16) - base
        n = int(r[cols["# Samples"]] or 0)
        ex = int(r[cols["Instructions Executed"]] or 0)
        f, line = where.get(off, ("?", 0))
        if f not in ranges:
            ranges[f] = function_ranges(f)
        fn = "?"
        for first, name in ranges[f]:
            if first <= line:
                fn = name
            else:
                break
        key = os.path.basename(f)
        by_file[key] += n
        by_func[(key, fn)] += n
        inst_by_func[(key, fn)] += ex
        by_line[(key, line)] += n
        total += n
    print(f"{total} samples, {len(inst)} SASS instructions")
    print("-- by file")
    for k, v in by_file.most_common():
        print(f"{100.0 * v / total:6.1f} %  {k}")
    print("-- by function (samples %, warp-instructions executed)")
    for k, v in by_func.most_common(30):
        print(f"{100.0 * v / total:6.1f} %  {inst_by_func[k]:>10}  {k[0]}:{k[1]}")
    # where each kind of stall sits: the per-instruction stall columns of the source page, by line
    for col in ("stall_long_sb", "stall_no_inst", "stall_wait", "stall_short_sb"):
        if col not in cols:
            continue
        by = collections.Counter()
        for r in inst:
            by[where.get(int(r[0], 16) - base, ("?", 0))] += int(r[cols[col]] or 0)
        tot = sum(by.values())
        if not tot:
            continue
        print(f"-- {col}: {tot} samples ({100.0 * tot / max(total, 1):.1f} % of all); top lines")
        for (f, ln), n in by.most_common(8):
            print(f"  {100.0 * n / tot:5.1f} %  {os.path.basename(f)}:{ln}")
    print("-- hottest lines")
    for k, v in by_line.most_common(25):
        print(f"{100.0 * v / total:6.1f} %  {k[0]}:{k[1]}")


if __name__ == "__main__":
    main()
