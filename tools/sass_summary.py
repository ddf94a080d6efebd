#!/usr/bin/env python
"""Per-kernel opcode histogram of numpyro_b200/libb200mc.so (cuobjdump -sass) ->
profiles/sass_summary.txt.  Shows which kernels are Blackwell-native: UTCHMMA (tcgen05.mma),
LDTM (tcgen05.ld), UTCBAR (tcgen05.commit), UBLKCP (cp.async.bulk), SYNCS (mbarrier), FFMA2
(packed fp32 FMA), MUFU.* (special-function unit)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "numpyro_b200", "libb200mc.so")
WATCH = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTCATOM", "UBLKCP", "UTMALDG", "SYNCS", "FFMA2", "FFMA",
         "FADD2", "FMUL2", "MUFU.EX2", "MUFU.LG2", "MUFU.RCP", "MUFU.RSQ", "MUFU.TANH", "LDS.128", "LDS",
         "LDG", "STG", "LDL", "STL", "SHFL", "BAR", "RED", "ATOM", "MEMBAR", "CALL"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(.*", "", name)
            cur = kernels.setdefault(name, collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur is not None:
            cur["_total"] += 1
            op = m.group(1)
            for w in WATCH:
                if op == w or op.startswith(w + "."):
                    cur[w] += 1
    head = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True,
                          text=True).stdout.strip()
    lines = [f"# cuobjdump -sass numpyro_b200/libb200mc.so  (sm_100a, tree at {head} + working copy)",
             "# Tile movement is 1-D cp.async.bulk (UBLKCP) into pre-tiled / pre-swizzled images, so there",
             "# is no UTMALDG (tensor-map TMA) by design; tcgen05 = UTCHMMA + LDTM + UTCBAR.", ""]
    tot = collections.Counter()
    for name, c in kernels.items():
        parts = [f"{w}={c[w]}" for w in WATCH if c[w]]
        lines.append(f"{name}: {c['_total']} instr; " + " ".join(parts))
        tot.update(c)
    lines += ["", "TOTAL: " + " ".join(f"{w}={tot[w]}" for w in WATCH if tot[w]),
              f"kernels: {len(kernels)}"]
    path = os.path.join(ROOT, "profiles", "sass_summary.txt")
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    print("\n".join(lines[-2:]))


if __name__ == "__main__":
    sys.exit(main())
