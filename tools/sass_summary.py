#!/usr/bin/env python
"""Opcode summary of the shipped library's SASS, per kernel: the mnemonics that show what the
kernels are made of - DPX (VIADDMNMX), texture fetches (TEX), surface stores (SUST), bulk async copies (UBLKCP),
mbarrier (SYNCS), warp reductions (REDUX), shared-memory loads (LDS), conversions (F2I) - and that there is no tensor-core
or TMA-tensor instruction (UTC*MMA / UTMALDG): the path has no dense contraction.

  python tools/sass_summary.py > profiles/r02_sass_opcodes.txt
"""
import collections, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tygrit_b200", "lib", "libtygrit_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
WATCH = ("VIADDMNMX", "VIMNMX", "TEX", "SUST", "UBLKCP", "SYNCS", "REDUX", "LDS", "STS", "F2I", "FFMA", "MUFU", "ATOMS", "ATOMG", "RED",
         "UTMALDG", "UTCHMMA", "UTCQMMA", "HMMA", "BAR", "SHFL", "VOTE", "STL", "LDL")
per = collections.OrderedDict()
name = None
arch = set(re.findall(r"arch = (sm_\w+)", out))
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        per.setdefault(name, collections.Counter())
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and name:
        per[name]["_total"] += 1
        op = m.group(1)
        for w in WATCH:
            if op.startswith(w):
                per[name][w] += 1
                break
print(f"# {os.path.relpath(lib, ROOT)}: arch {sorted(arch)}; instructions per kernel and the watched opcode families")
tot = collections.Counter()
for k, c in per.items():
    if c["_total"] < 40:
        continue
    tot.update(c)
    print(f"{k[:60]:60s} n={c['_total']:6d}  " + "  ".join(f"{w}={c[w]}" for w in WATCH if c[w]))
print("TOTAL".ljust(60) + f" n={tot['_total']:6d}  " + "  ".join(f"{w}={tot[w]}" for w in WATCH))
