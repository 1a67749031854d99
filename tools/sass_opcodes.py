#!/usr/bin/env python
"""Per-kernel SASS opcode histogram of libb200xp.so (cuobjdump -sass), so that the claims in DESIGN.md are
checkable without a toolchain: UTMALDG / UTMASTG / SYNCS (TMA + mbarrier) in gae_tma_kernel, LDG.E.*.256
(256-bit loads) in per_sample_kernel, LDGSTS (cp.async) in the token GAE, SHFL / VOTE / MATCH in the warp-level
tree set, ld/st on peer-mapped (system-scope) addresses in the peer kernels.
usage: python tools/sass_opcodes.py [lib] > profiles/rNN_sass_opcodes.txt"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "agilerl_b200/_lib/libb200xp.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
KEY = re.compile(r"UTMA|UBLKCP|SYNCS|LDGSTS|\.256|\.128|SHFL|VOTE|MATCH|ATOM|RED\.|DADD|DMUL|DFMA|F2F|\.SYS|BAR|CCTL|MEMBAR|ACQBULK|UTC|HMMA|LDTM")
arch = None
fn = None
hist: dict = {}
for ln in out.splitlines():
    m = re.search(r"arch = (sm_\w+)", ln)
    if m:
        arch = m.group(1)
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
        fn = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", fn)
        fn = re.sub(r"\(.*", "", fn)
        hist[fn] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", ln)
    if m and fn:
        hist[fn][m.group(1)] += 1
print(f"# {lib}: {len(hist)} kernels, arch {arch} only; opcode counts per kernel (static SASS)")
for fn in sorted(hist):
    h = hist[fn]
    total = sum(h.values())
    keys = {k: v for k, v in h.items() if KEY.search(k)}
    top = ", ".join(f"{k} {v}" for k, v in h.most_common(6))
    print(f"\n{fn}\n  instructions {total}; top: {top}")
    if keys:
        print("  notable: " + ", ".join(f"{k} {v}" for k, v in sorted(keys.items())))
