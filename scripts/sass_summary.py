"""Counts, per kernel of libexg.so, the SASS mnemonics that prove which hardware path the kernel uses (B200_PROFILING.md):
UTCHMMA / UTCQMMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UTMALDG / UTMASTG / UTMAREDG (TMA load / store /
reduce-add), UTCBAR (tcgen05.commit), SYNCS (mbarrier), HMMA (legacy mma.sync), plus registers and spills from the
cubin's resource usage.  Runs anywhere cuobjdump is installed (no GPU needed):

    python scripts/sass_summary.py > profiles/sass_r02_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "exoground_b200", "libexg.so")
KEYS = ("UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "SYNCS", "HMMA",
        "FFMA", "MUFU", "LDG", "STG", "LDS", "STS", "ATOM", "RED")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m:
            op = m.group(1)
            counts[cur]["_total"] += 1
            for k in KEYS:
                if op == k or op.startswith(k):
                    counts[cur][k] += 1
                    break
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    usage, fn = {}, None
    for line in res.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            fn = m.group(1)
            continue
        if fn and "REG:" in line:
            usage[fn] = dict(re.findall(r"(REG|STACK|SHARED|LOCAL):(\d+)", line))
            fn = None
    names = demangle(list(counts))
    print(f"# SASS summary of {os.path.relpath(LIB, ROOT)} (cuobjdump -sass / -res-usage, sm_100a), one line per kernel")
    print("# columns: instructions | tcgen05.mma (UTCHMMA) | tcgen05.commit (UTCBAR) | tcgen05.ld (LDTM) | TMA load / store /"
          " reduce (UTMALDG/UTMASTG/UTMAREDG) | mbarrier (SYNCS) | legacy HMMA | regs | stack bytes")
    tc = []
    for fn, c in counts.items():
        u = usage.get(fn, {})
        short = names.get(fn, fn).replace("(anonymous namespace)::", "")
        short = re.sub(r"^void ", "", re.sub(r"\(.*", "", short))
        row = (f"{short[:86]:86s} {c['_total']:6d} | mma {c['UTCHMMA'] + c['UTCQMMA']:3d} | commit {c['UTCBAR']:3d} | "
               f"ldtm {c['LDTM']:3d} | tma {c['UTMALDG']:3d}/{c['UTMASTG']:3d}/{c['UTMAREDG']:3d} | syncs {c['SYNCS']:3d} | "
               f"hmma {c['HMMA']:3d} | regs {u.get('REG', '?'):>3s} | stack {u.get('STACK', '?'):>4s}")
        print(row)
        if c["UTCHMMA"] + c["UTCQMMA"]:
            tc.append(short)
    print(f"# {len(counts)} kernels; {len(tc)} issue tcgen05.mma; "
          f"{sum(1 for c in counts.values() if c['UTMALDG'] or c['UTMASTG'] or c['UTMAREDG'])} use TMA; "
          f"{sum(1 for c in counts.values() if c['HMMA'])} use legacy HMMA")


if __name__ == "__main__":
    sys.exit(main())
