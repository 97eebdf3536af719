#!/usr/bin/env python3
"""Per-source-line instruction / stall-sample shares of the first kernel in an .ncu-rep (needs -lineinfo + --import-source on).
   python tools/ncu_lines.py <rep> [N]"""
import csv, subprocess, sys
args = [a for a in sys.argv[1:] if not a.startswith("--")]; rep = args[0]; N = int(args[1]) if len(args) > 1 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
h = None; fpath = ""; out = []; kernels = 0
for r in rows:
    if r and r[0] == "File Path": fpath = r[1].split("/")[-1]; continue
    if r and r[0] == "Function Name":
        kernels += 1; continue
    if r and r[0] == "Line No": h = r; continue
    if h is None or len(r) != len(h): continue
    if r[0] == "": continue          # SASS rows under a source line
    iI = h.index("Instructions Executed"); iS = h.index("# Samples")
    try: out.append((int(r[iI]), int(r[iS]), fpath, r[0], r[1][:120]))
    except ValueError: pass
tot = sum(o[0] for o in out) or 1; tots = sum(o[1] for o in out) or 1
out.sort(key=(lambda o: (o[1], o[0])) if "--by-samples" in sys.argv else None, reverse=True)
print("total warp-instructions %d, samples %d" % (tot, tots))
for n, s, f, l, src in out[:N]:
    print("%5.1f%% %5.1f%%  %s:%s  %s" % (100 * n / tot, 100 * s / tots, f, l, src.strip()))
