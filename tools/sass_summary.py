"""Per-kernel tcgen05 / TMEM / TMA evidence from the shipped library:

    python tools/sass_summary.py > profiles/r2_sass_summary.md

Runs ``cuobjdump -sass`` on ``maestro_b200/libmaestro_b200.so`` and counts, per kernel, the SASS mnemonics that prove
the Blackwell data path (B200_PROFILING.md): UTCHMMA (tcgen05.mma, ``.2CTA`` = cta_group::2), LDTM (tcgen05.ld),
UTMALDG / UTMASTG / UTMAPF (TMA tensor load / store / L2 prefetch), UBLKCP (cp.async.bulk), FFMA2 / FADD2 / FMUL2
(packed fp32), MUFU.EX2, SYNCS (mbarrier).
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "maestro_b200", "libmaestro_b200.so")
PATTERNS = ["UTCHMMA.2CTA", "UTCHMMA", "LDTM", "UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "FFMA2", "FADD2", "FMUL2",
            "MUFU.EX2", "SYNCS"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    counts, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        counts[cur]["total"] += 1
        for p in PATTERNS:
            if op.startswith(p):
                counts[cur][p] += 1
                break
    names = demangle(list(counts))
    cols = [p for p in PATTERNS if any(c[p] for c in counts.values())]
    print(f"# SASS summary of maestro_b200/libmaestro_b200.so ({len(counts)} kernels, sm_100a)\n")
    print("| kernel | instr | " + " | ".join(cols) + " |")
    print("|---|---:|" + "---:|" * len(cols))
    tot = collections.Counter()
    for k, c in sorted(counts.items(), key=lambda kv: names[kv[0]]):
        short = names[k].replace("(anonymous namespace)::", "").replace("void ", "")
        short = re.sub(r"\(.*", "", short)
        if not any(c[p] for p in cols):
            continue
        print(f"| `{short}` | {c['total']} | " + " | ".join(str(c[p]) if c[p] else "" for p in cols) + " |")
        tot.update(c)
    print(f"| **all listed** | {tot['total']} | " + " | ".join(str(tot[p]) for p in cols) + " |")


if __name__ == "__main__":
    sys.exit(main())
