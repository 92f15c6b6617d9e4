"""Counts the tensor-core / TMA / TMEM SASS mnemonics per kernel of the built library (profiles/r02_sass_summary.csv).
Run here (no GPU): python scripts/sass_summary.py > profiles/r02_sass_summary.csv"""
import collections
import re
import subprocess
import sys

LIB = "gnn_b200/lib/libgnn_b200.so"
COLS = ["UTCIMMA", "UTCHMMA", "UTCOMMA", "UTMALDG", "LDTM", "STTM", "UTCBAR", "HMMA", "FFMA2", "LDG.E.128", "LDGMC"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    names = {}
    counts = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        for c in COLS:
            if re.search(r"\b" + re.escape(c), line):
                counts[cur][c] += 1
    mangled = list(counts)
    dem = subprocess.run(["c++filt"], input="\n".join(mangled), capture_output=True, text=True).stdout.splitlines()
    for a, b in zip(mangled, dem):
        b = b.replace("(anonymous namespace)::", "").replace("void ", "")
        names[a] = re.sub(r"\((?!anonymous).*$", "", b)
    print("# cuobjdump -sass " + LIB + ", occurrences per kernel (UTCIMMA/UTCHMMA/UTCOMMA = tcgen05.mma kind::i8 / "
          "kind::f16 / kind::mxf4 block-scaled, UTMALDG = TMA load, LDTM/STTM = tcgen05.ld/st, HMMA = legacy mma.sync, FFMA2 = packed fp32 fma (fma.rn.f32x2), "
          "LDGMC = multimem.ld_reduce)")
    print("kernel," + ",".join(COLS))
    for k in sorted(counts, key=lambda k: names[k]):
        print('"' + names[k] + '",' + ",".join(str(counts[k][c]) for c in COLS))


if __name__ == "__main__":
    sys.exit(main())
