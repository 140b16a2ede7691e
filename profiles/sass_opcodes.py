"""Per-kernel counts of the Blackwell-only SASS opcodes in the built library:

    python profiles/sass_opcodes.py > profiles/sass_opcodes.md        (needs cuobjdump; no GPU)

UTCHMMA / UTCQMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st (TMEM), UTMALDG / UTMASTG / UTMAREDG = TMA tensor load /
store / reduce-add, UBLKCP = bulk (1-D) TMA copy, SYNCS = mbarrier, UTCBAR = tcgen05.commit; FFMA2 etc. are listed to show
the packed-pair math.  BRA.U.ANY = the back edge of a waterfall loop ptxas wraps around an instruction whose uniform-register
operands it cannot prove warp-uniform (0 for the hot kernels since round 2d, profiles/r2d_issue_path.md); ELECT = elect.sync.
See /opt/skills/guides/B200_PROFILING.md for the mnemonics."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "acai_omr_b200", "libacai_b200.so")
OPS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "UTCBAR", "SYNCS", "FFMA2", "HMMA", "MUFU.EX2", "ELECT",
       "BRA.U.ANY"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
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
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            op = m.group(1)
            for key in OPS:
                if op == key or op.startswith(key + "."):
                    counts[cur][key] += 1
            counts[cur]["_total"] += 1
    names = demangle(list(counts))
    rows = []
    for k, c in counts.items():
        short = re.sub(r"\(.*", "", names[k]).replace("acai::", "").replace("(anonymous namespace)::", "")
        short = re.sub(r"^void ", "", short)
        rows.append((short, c))
    rows.sort(key=lambda r: (-(r[1]["UTCHMMA"] + r[1]["UTMALDG"] + r[1]["UBLKCP"] + r[1]["LDTM"]), r[0]))
    tot = collections.Counter()
    for _, c in rows:
        tot.update(c)
    print("# SASS opcode summary of `acai_omr_b200/libacai_b200.so` (sm_100a)\n")
    print(f"`cuobjdump -sass` over {len(rows)} kernels, {tot['_total']} instructions.  Totals: "
          + ", ".join(f"{k} {tot[k]}" for k in OPS if tot[k]) + ".\n")
    print("Kernels that use tcgen05 / TMEM / TMA (one row per template instantiation):\n")
    print("| kernel | instr | " + " | ".join(OPS) + " |")
    print("|---|---|" + "---|" * len(OPS))
    for short, c in rows:
        if not any(c[k] for k in OPS[:9]):
            continue
        print(f"| `{short[:110]}` | {c['_total']} | " + " | ".join(str(c[k]) if c[k] else "" for k in OPS) + " |")
    rest = [s for s, c in rows if not any(c[k] for k in OPS[:9])]
    print(f"\nThe other {len(rest)} kernels are SIMT (row-wise, fp32-exact and glue kernels): "
          + ", ".join(sorted({re.sub(r'<.*', '', s) for s in rest})) + ".")


if __name__ == "__main__":
    sys.exit(main())
