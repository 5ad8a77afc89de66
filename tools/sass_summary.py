"""Per-kernel counts of the Blackwell tensor-core / TMA instructions in the built library.

  python tools/sass_summary.py > profiles/sass_summary.txt

Runs `cuobjdump -sass` on exorl_b200/lib/libexorl_b200.so and counts, per kernel, the SASS
mnemonics that prove the tcgen05 / TMEM / TMA path (B200_PROFILING.md): UTCHMMA (tcgen05.mma),
LDTM (tcgen05.ld from TMEM), UTMALDG (TMA tensor load), UTCBAR (tcgen05.commit), SYNCS (mbarrier),
UBLKCP (bulk copy), plus HMMA/FFMA for contrast.  Evidence file, not product code.
"""
import collections
import pathlib
import re
import subprocess
import sys

ROOT = pathlib.Path(__file__).resolve().parent.parent
LIB = ROOT / "exorl_b200" / "lib" / "libexorl_b200.so"
OPS = ("UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "SYNCS", "UBLKCP",
       "HMMA", "FFMA", "FFMA2", "LDG", "STG")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True,
                          check=True).stdout
    demangle = {}
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
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            op = m.group(1)
            counts[cur][op] += 1
            counts[cur]["_total"] += 1
    names = list(counts)
    filt = subprocess.run(["c++filt", "-p"], input="\n".join(names), capture_output=True,
                          text=True).stdout.splitlines()
    demangle = dict(zip(names, filt))
    print(f"# cuobjdump -sass {LIB.relative_to(ROOT)}  ({len(names)} kernels, sm_100a)")
    print(f"# {'kernel':<58} {'instr':>7} " + " ".join(f"{o:>7}" for o in OPS))
    tot = collections.Counter()
    for n in names:
        c = counts[n]
        short = demangle.get(n, n)
        short = re.sub(r"^exb::", "", short)
        print(f"{short[:60]:<60} {c['_total']:>7} " + " ".join(f"{c[o]:>7}" for o in OPS))
        tot.update(c)
    print(f"{'TOTAL':<60} {tot['_total']:>7} " + " ".join(f"{tot[o]:>7}" for o in OPS))


if __name__ == "__main__":
    sys.exit(main())
