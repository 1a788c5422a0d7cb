"""SASS evidence: per kernel of libgmm_b200.so, how many tcgen05 / TMEM / bulk-copy instructions it contains.
python profiles/sass_summary.py > profiles/sass_summary_r02.txt   (cuobjdump only; no GPU needed)

Mnemonics (B200_PROFILING.md): UTCHMMA = tcgen05.mma kind::f16, UTCQMMA = kind::f8f6f4, UTCIMMA = kind::i8,
UTCBAR = tcgen05.commit, LDTM = tcgen05.ld, UBLKCP = cp.async.bulk (1-D bulk copy, TMA engine), SYNCS = mbarrier,
MUFU.EX2 = ex2.approx, VIMNMX3 / FMNMX3 = three-input integer / float min-max, IDP.4A = dp4a."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "rasr_b200", "libgmm_b200.so")
KEYS = ["UTCHMMA", "UTCQMMA", "UTCIMMA", "UTCBAR", "LDTM", "UBLKCP", "SYNCS", "MUFU.EX2", "FMNMX3", "FMNMX", "VIMNMX3", "IMAD",
        "LEA", "IDP.4A", "STG.E.128", "STG", "STS.128", "LDS.128", "STL", "LDL"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    fn, counts, size = None, {}, {}
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            fn = m.group(1)
            counts[fn] = collections.Counter()
            size[fn] = 0
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if fn and m:
            op = m.group(1)
            size[fn] += 1
            for k in KEYS:
                if op == k or op.startswith(k + ".") or (k in ("STG.E.128", "STS.128", "LDS.128", "MUFU.EX2") and op.startswith(k)):
                    counts[fn][k] += 1
    names = {}
    for f in counts:
        d = subprocess.run(["c++filt", f], capture_output=True, text=True).stdout.strip()
        d = re.sub(r"\(anonymous namespace\)::", "", d)
        d = re.sub(r"\(.*$", "", d)
        d = re.sub(r"^void ", "", d)
        names[f] = d
    print("libgmm_b200.so, sm_100a, static instruction counts per kernel (cuobjdump -sass)")
    print(f"{'kernel':44s} {'instr':>6s} " + " ".join(f"{k[:8]:>8s}" for k in KEYS))
    for f in sorted(counts, key=lambda f: names[f]):
        c = counts[f]
        if not any(c[k] for k in ("UTCHMMA", "UTCQMMA", "UTCIMMA", "LDTM", "UBLKCP", "IDP.4A")):
            continue
        print(f"{names[f][:44]:44s} {size[f]:6d} " + " ".join(f"{c[k]:8d}" for k in KEYS))


if __name__ == "__main__":
    sys.exit(main())
