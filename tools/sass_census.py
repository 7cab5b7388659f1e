"""Per-kernel instruction census of the built library: how many tcgen05 / TMEM / TMA / mma.sync SASS instructions each
kernel contains (cuobjdump -sass kair_b200/libkair_b200.so).  Writes profiles/sass_census_r2.txt.

    UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / .st (TMEM), UTMALDG / UTMASTG / UTMAREDG = TMA load / store /
    reduce, HMMA = mma.sync (legacy tensor-core path), SYNCS = mbarrier ops, UTCBAR = tcgen05.commit."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "kair_b200", "libkair_b200.so")
PATTERNS = ["UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "UTCBAR", "HMMA", "SYNCS"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts, total, cur = collections.OrderedDict(), {}, None
    for line in sass.split("\n"):
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            total[cur] = 0
            continue
        if cur is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            total[cur] += 1
            op = m.group(1).split(".")[0]
            for p in PATTERNS:
                if op.startswith(p):
                    counts[cur][p] += 1
    names = demangle(list(counts))
    path = os.path.join(ROOT, "profiles", "sass_census_r2.txt")
    with open(path, "w") as f:
        f.write("# SASS instruction census of kair_b200/libkair_b200.so (sm_100a), tools/sass_census.py\n")
        f.write("# UTCHMMA tcgen05.mma | LDTM/STTM tcgen05.ld/st | UTMALDG/UTMASTG/UTMAREDG TMA load/store/reduce | "
                "HMMA mma.sync | SYNCS mbarrier\n")
        f.write("%-88s %7s " % ("kernel", "instrs") + " ".join("%8s" % p for p in PATTERNS) + "\n")
        for k, c in counts.items():
            short = re.sub(r"\(anonymous namespace\)::", "", names.get(k, k))
            short = re.sub(r"\(.*\)$", "", short)[:88]
            f.write("%-88s %7d " % (short, total[k]) + " ".join("%8d" % c[p] for p in PATTERNS) + "\n")
    print("wrote", path, "(%d kernels)" % len(counts))


if __name__ == "__main__":
    sys.exit(main())
