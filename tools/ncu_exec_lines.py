"""Executed warp instructions per source line (and per opcode) of an .ncu-rep, de-duplicated by SASS address.
usage: ncu_exec_lines.py rep [n] [units]  - units: divide counts by this many work items (e.g. point-evaluations / 32)"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
units = float(sys.argv[3]) if len(sys.argv) > 3 else None
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
hdr = cur = line = None
seen = {}
text = {}
for r in csv.reader(io.StringIO(src)):
    if not r: continue
    if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r[0] == 'Line No': hdr = r; continue
    if r[0] == 'Function Name' or hdr is None: continue
    if r[0] != '': line = (cur, int(r[0])); text[line] = r[1].strip()[:90]; continue
    if r[2] in ('...', '-') or r[2] in seen: continue
    try: seen[r[2]] = (int(r[6]), int(r[7]), r[3].strip()[:64], line)
    except ValueError: continue
ex = sum(v[1] for v in seen.values())
byline, byop = collections.Counter(), collections.Counter()
for s, e, t, l in seen.values():
    byline[l] += e
    w = t.split(); o = (w[1] if w[0].startswith('@') else w[0]).split('.')[0]
    byop[o] += e
sc = (lambda v: f"{v / units:8.2f}/unit") if units else (lambda v: "")
print(f"warp instructions {ex} {sc(ex)}")
for k, v in byop.most_common(25): print(f"  {k:12s} {100 * v / ex:5.1f}% {sc(v)}")
for l, v in byline.most_common(n): print(f"{100 * v / ex:5.1f}% {sc(v)}  {l[0]}:{l[1]}  {text.get(l, '')}")
