#!/usr/bin/env python
"""SASS instruction count per source line of one kernel (no GPU needed): where the code size of a variant sits.
Usage: python tools/sass_lines.py <lib.so> <kernel-name-substring> [file-substring]"""
import collections, os, re, subprocess, sys, tempfile

lib, kern = os.path.abspath(sys.argv[1]), sys.argv[2]
only = sys.argv[3] if len(sys.argv) > 3 else ""
with tempfile.TemporaryDirectory() as td:
    subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=td, check=True, capture_output=True)
    cubin = [f for f in os.listdir(td) if f.endswith(".cubin")][0]
    out = subprocess.run(["nvdisasm", "--print-line-info", cubin], cwd=td, capture_output=True, text=True).stdout
cur_fn, cur_line, cnt, total = None, None, collections.Counter(), 0
for l in out.splitlines():
    m = re.match(r"\.text\.(\S+):", l)
    if m:
        cur_fn = m.group(1); cur_line = None; continue
    if cur_fn is None or kern not in cur_fn: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur_line = (m.group(1).split("/")[-1], int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4}\*/", l):
        total += 1
        if cur_line: cnt[cur_line] += 1
byfile = collections.Counter()
for (f, _), v in cnt.items(): byfile[f] += v
print(f"{kern}: {total} instructions;", dict(byfile))
for (f, ln), v in sorted(cnt.items()):
    if only in f: print(f"{f}:{ln}  {v}")
