#!/usr/bin/env python
"""Per-kernel counts of the Blackwell-native SASS mnemonics in lib/libb200llama.so
(cuobjdump -sass): UTCHMMA (tcgen05.mma), UTMALDG (TMA tensor loads), LDTM (tcgen05.ld),
UTCBAR (tcgen05.commit), UBLKCP (cp.async.bulk 1-D), UBLKPF (bulk L2 prefetch), HMMA (mma.sync),
SYNCS (mbarrier), plus registers from `cuobjdump -res-usage`.

    python tools/sass_counts.py profiles/r02_sass_counts.md
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "llm-inference_b200", "lib", "libb200llama.so")
KEYS = ["UTCHMMA", "UTCHMMA.2CTA", "UTMALDG", "LDTM", "STTM", "UTCBAR", "UBLKCP", "UBLKPF", "HMMA", "SYNCS", "FFMA/HFMA2.MMA/FHFMA"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def main(dst):
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
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        c = counts[cur]
        c["_total"] += 1
        if op.startswith("UTCHMMA"):
            c["UTCHMMA"] += 1
            if ".2CTA" in op:
                c["UTCHMMA.2CTA"] += 1
        elif op.startswith("UTMALDG"):
            c["UTMALDG"] += 1
        elif op.startswith("LDTM"):
            c["LDTM"] += 1
        elif op.startswith("STTM"):
            c["STTM"] += 1
        elif op.startswith("UTCBAR"):
            c["UTCBAR"] += 1
        elif op.startswith("UBLKCP"):
            c["UBLKCP"] += 1
        elif op.startswith("UBLKPF"):
            c["UBLKPF"] += 1
        elif op.startswith("HMMA"):
            c["HMMA"] += 1
        elif op.startswith("SYNCS"):
            c["SYNCS"] += 1
        elif op.startswith("FHFMA") or op.startswith("FFMA") or op.startswith("HFMA2"):
            c["FFMA/HFMA2.MMA/FHFMA"] += 1
    names = demangle(list(counts))
    rows = []
    tot = collections.Counter()
    for k, c in counts.items():
        tot.update(c)
        if any(c[x] for x in KEYS[:9]):
            nm = re.sub(r"\(.*", "", names[k]).replace("void ", "").replace("b200::", "")
            rows.append((nm, c))
    with open(dst, "w") as f:
        f.write("SASS mnemonic counts per kernel, `cuobjdump -sass llm-inference_b200/lib/libb200llama.so` "
                "(sm_100a; built by `__graft_entry__.build()`), produced by `tools/sass_counts.py`.\n"
                "Kernels without any tensor-core / TMA / bulk-copy / mbarrier instruction are left out.\n\n")
        f.write("| kernel | " + " | ".join(KEYS) + " | instructions |\n|---|" + "---|" * (len(KEYS) + 1) + "\n")
        for nm, c in sorted(rows, key=lambda r: r[0]):
            f.write(f"| `{nm}` | " + " | ".join(str(c[k]) for k in KEYS) + f" | {c['_total']} |\n")
        f.write("\n**Whole library:** " + ", ".join(f"{k} {tot[k]}" for k in KEYS) + f"; {len(counts)} kernels, "
                f"{tot['_total']} instructions.\n")
    print("wrote", dst, "-", ", ".join(f"{k} {tot[k]}" for k in KEYS[:9]))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_sass_counts.md"))
