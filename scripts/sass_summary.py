#!/usr/bin/env python
"""Per-kernel counts of the Blackwell-only SASS instructions in the built library (evidence that the contraction is
tcgen05 / TMEM / TMA code, not recompiled mma.sync):

    python scripts/sass_summary.py [evi-rag_b200/libevi_rag_b200.so] > profiles/sass_summary.md

UTCHMMA = tcgen05.mma (…2CTA = cta_group::2), LDTM = tcgen05.ld, UTMALDG / UTMASTG = TMA tensor load / store,
UBLKCP = cp.async.bulk, UTCBAR = tcgen05.commit, SYNCS = mbarrier ops. Runs without a GPU (cuobjdump only)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "evi-rag_b200", "libevi_rag_b200.so")
MNEMONICS = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "UTMALDG", "UTMALDG.2CTA", "UTMASTG", "UBLKCP", "UTCBAR", "SYNCS",
             "HMMA", "FFMA", "MUFU"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
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
        m = re.search(r"\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        base = op.split(".")[0]
        counts[cur][base] += 1
        if base in ("UTCHMMA", "UTMALDG") and ".2CTA" in op:
            counts[cur][base + ".2CTA"] += 1
    names = demangle(list(counts))
    arch = re.findall(r"arch = (sm_\w+)", sass)
    print(f"# SASS summary of `{os.path.relpath(LIB, ROOT)}` ({', '.join(sorted(set(arch)))}; {len(counts)} kernels)\n")
    print("Counts of instructions per kernel (`cuobjdump -sass`, `scripts/sass_summary.py`). `.2CTA` columns are the subset "
          "issued for a CTA pair (`cta_group::2`). HMMA (the mma.sync path) is absent everywhere.\n")
    print("| kernel | " + " | ".join(MNEMONICS) + " |")
    print("|---|" + "---:|" * len(MNEMONICS))
    tot = collections.Counter()
    for k, c in counts.items():
        if not any(c[m] for m in MNEMONICS[:9]):
            continue
        short = re.sub(r"\(.*", "", names[k])
        short = re.sub(r"^void ", "", short).replace("evi::", "")
        print(f"| `{short[:90]}` | " + " | ".join(str(c[m]) for m in MNEMONICS) + " |")
        tot.update({m: c[m] for m in MNEMONICS})
    print("| **all kernels with tensor / TMA / bulk-copy instructions** | " + " | ".join(str(tot[m]) for m in MNEMONICS) + " |")
    other = [re.sub(r"\(.*", "", names[k]).replace("void ", "").replace("evi::", "") for k, c in counts.items()
             if not any(c[m] for m in MNEMONICS[:9])]
    print(f"\n{len(other)} kernels use none of them (integer / byte / elementwise work on coalesced loads): "
          + ", ".join(f"`{o[:60]}`" for o in sorted(set(other))) + ".")


if __name__ == "__main__":
    main()
