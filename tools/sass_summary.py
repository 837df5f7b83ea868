#!/usr/bin/env python
"""Per-kernel counts of the Blackwell-specific SASS instructions in libdlco_b200.so (cuobjdump -sass):
tcgen05 MMAs (UTC*MMA), TMEM loads / stores (LDTM / STTM), TMA tensor and bulk copies (UTMALDG / UBLKCP /
UTMASTG), mbarrier ops (SYNCS), and the register / static shared memory footprint.  Evidence that the tensor
paths are tcgen05 / TMEM / TMA kernels, written to profiles/sass_summary.txt:

    python tools/sass_summary.py > profiles/sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "opencv-dlco_b200", "libdlco_b200.so")
PATS = [("UTC*MMA", r"\bUTC[A-Z0-9]*MMA"), ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"), ("UTMALDG", r"\bUTMALDG"),
        ("UTMASTG", r"\bUTMASTG"), ("UBLKCP", r"\bUBLKCP"), ("SYNCS", r"\bSYNCS"), ("HMMA/IMMA (legacy)", r"\b[HI]MMA"),
        ("DFMA/DMUL/DADD", r"\bD(FMA|MUL|ADD)\b"), ("FFMA", r"\bFFMA")]


def demangle(names):
    try:
        out = subprocess.run(["c++filt"], input="\n".join(names), stdout=subprocess.PIPE, text=True).stdout.splitlines()
        return dict(zip(names, out))
    except OSError:
        return {n: n for n in names}


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else LIB
    sass = subprocess.run(["cuobjdump", "-sass", lib], stdout=subprocess.PIPE, text=True, check=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", lib], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True).stdout
    usage = {}
    cur = None
    for ln in res.splitlines():
        m = re.match(r"\s*Function (\S+):", ln)
        if m:
            cur = m.group(1)
            continue
        m = re.search(r"REG:(\d+).*?SHARED:(\d+)", ln)
        if m and cur:
            usage[cur] = (int(m.group(1)), int(m.group(2)))
    counts = collections.OrderedDict()
    cur = None
    arch = ""
    for ln in sass.splitlines():
        m = re.match(r"\s*arch = (\S+)", ln)
        if m:
            arch = m.group(1)
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            counts[cur]["arch"] = arch
            continue
        if cur and re.match(r"\s+/\*[0-9a-f]{4}\*/", ln):
            counts[cur]["total"] += 1
            for name, pat in PATS:
                if re.search(pat, ln):
                    counts[cur][name] += 1
    names = demangle(list(counts))
    print("# %s  (cuobjdump -sass; one line per kernel)" % os.path.relpath(lib, ROOT))
    cols = [n for n, _ in PATS]
    print("%-64s %-7s %5s %6s %7s  %s" % ("kernel", "arch", "regs", "smem", "instrs", "  ".join(cols)))
    for fn, c in counts.items():
        short = names.get(fn, fn).replace("(anonymous namespace)::", "").replace("<unnamed>::", "").replace("void ", "")
        short = re.sub(r"\(.*$", "", short)
        reg, sh = usage.get(fn, (0, 0))
        print("%-64s %-7s %5d %6d %7d  %s" % (short[:64], c["arch"], reg, sh, c["total"],
                                             "  ".join("%*d" % (len(n), c[n]) for n in cols)))


if __name__ == "__main__":
    main()
