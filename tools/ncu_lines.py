#!/usr/bin/env python3
"""Dev helper: join an ncu report's per-SASS-instruction counts with source lines (nvdisasm -g).

    python tools/ncu_lines.py report.ncu-rep kernel_substring cubin_glob_dir
"""
import csv, io, re, subprocess, sys, glob, collections, os, tempfile
rep, kname = sys.argv[1], sys.argv[2]
lib = os.path.join(os.path.dirname(__file__), "..", "enph459-super-resolution_b200", "libsr459.so")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
line_of = {}
for f in glob.glob(tmp + "/*.cubin"):
    if os.path.basename(f).count("-") > 2: continue
    txt = subprocess.run(["nvdisasm", "-g", "-c", f], capture_output=True, text=True).stdout
    if kname not in txt: continue
    cur = None; infn = False
    for ln in txt.splitlines():
        if ln.startswith("//---") and ".text." in ln: infn = kname in ln
        if not infn: continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
        m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', ln)
        if m: line_of[int(m.group(1), 16)] = (cur, m.group(2))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
iA, iN, iS = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
per = collections.Counter(); samp = collections.Counter(); tot = 0; base = None; seen = set()
for r in rows[2:]:
    if len(r) < len(hdr):
        if r and r[0] == "Kernel Name" and seen: break
        continue
    try: addr = int(r[iA], 16); n = int(r[iN]); s = int(r[iS] or 0)
    except ValueError: continue
    if base is None: base = addr
    off = addr - base
    if off in seen: break
    seen.add(off)
    key = line_of.get(off, (None, ""))[0]
    per[key] += n; samp[key] += s; tot += n
srcs = {}
print("total warp instr", tot)
for key, n in per.most_common(40):
    text = ""
    if key:
        p = os.path.join(os.path.dirname(lib), "csrc", key[0])
        if p not in srcs and os.path.exists(p): srcs[p] = open(p).read().splitlines()
        if p in srcs and key[1] <= len(srcs[p]): text = srcs[p][key[1] - 1].strip()[:90]
    print(f"{n:9d} {100*n/tot:5.1f}%  stall-samples {samp[key]:5d}  {key}  {text}")
