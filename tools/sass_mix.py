#!/usr/bin/env python
"""SASS instruction mix of the k_render variants in libpt_b200.so (no GPU needed): instruction count, the mnemonics that
matter for this path (FP32 pipe, local memory, 16-byte global loads, warp collectives, IEEE slow-path calls) and the proof
that there is nothing tensor- or TMA-shaped in it.  usage: python tools/sass_mix.py [lib.so] > profiles/rNN_sass_mix.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "pathtracer_b200", "libpt_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
fn, mix = None, collections.defaultdict(collections.Counter)
for l in out.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_.]+)?)", l)
    if m and fn:
        op = m.group(1)
        mix[fn][op.split(".")[0]] += 1
        if op.startswith("LDG.E.128") or op.startswith("LDG.E.64"):
            mix[fn][op.rsplit(".CONSTANT", 1)[0]] += 0            # keep the key visible below
            mix[fn]["LDG.128" if ".128" in op else "LDG.64"] += 1
print(f"# {os.path.basename(lib)}: SASS instruction mix per kernel (cuobjdump -sass); columns are static instruction counts")
cols = ["total", "FFMA", "FMUL", "FADD", "FMNMX", "FMNMX3", "MUFU", "IMAD", "LOP3", "LDG", "LDG.128", "LDL", "STL", "LDS", "STS", "SHFL", "VOTE", "CREDUX", "CALL", "BRA", "BSSY"]
print(f"{'kernel':44s} " + " ".join(f"{c:>7s}" for c in cols))
for f in sorted(mix):
    if not ("k_render" in f or "k_resolve" in f or "k_trace" in f or "k_primary" in f or "pack" in f):
        continue
    c = mix[f]
    c["total"] = sum(v for k, v in c.items() if "." not in k)
    print(f"{f[:44]:44s} " + " ".join(f"{c.get(k, 0):7d}" for k in cols))
tensor = sorted({k for c in mix.values() for k in c if re.match(r"(UTC|UTMA|LDTM|STTM|HMMA|IMMA|QMMA|OMMA|UBLKCP|UTMALDG)", k)})
print(f"\ntensor-core / TMEM / TMA mnemonics in any kernel of the library: {tensor or 'none'} (by design: nothing on this path is a "
      "dense contraction or a bulk tile copy -- BASELINE.json north_star)")
