#!/usr/bin/env python
"""Per-function SASS opcode histogram: python tools/sass_hist.py <cubin|obj|exe> [name-filter] [top]"""
import collections
import re
import subprocess
import sys

out = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
flt = sys.argv[2] if len(sys.argv) > 2 else ""
top = int(sys.argv[3]) if len(sys.argv) > 3 else 12
fn, hist = None, None
funcs = []
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn, hist = m.group(1), collections.Counter()
        funcs.append((fn, hist))
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m and hist is not None:
        hist[m.group(1)] += 1
for fn, hist in funcs:
    if flt in fn:
        name = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip()[:110]
        print(f"{name}: {sum(hist.values())} instrs: " + ", ".join(f"{k} {v}" for k, v in hist.most_common(top)))
