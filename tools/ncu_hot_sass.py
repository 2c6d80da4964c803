"""SASS instructions of an .ncu-rep in address order with their executed counts, restricted to those executed at least
`frac` of the most-executed instruction (the steady-state loops).  usage: ncu_hot_sass.py rep [frac]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
hdr = cur = line = None
seen = {}
for r in csv.reader(io.StringIO(src)):
    if not r: continue
    if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r[0] == 'Line No': hdr = r; continue
    if r[0] == 'Function Name' or hdr is None: continue
    if r[0] != '': line = (cur, int(r[0])); continue
    if r[2] in ('...', '-') or r[2] in seen: continue
    try: seen[r[2]] = (int(r[7]), int(r[6]), r[3].strip()[:90], line)
    except ValueError: continue
mx = max(v[0] for v in seen.values())
tot = 0
for a in sorted(seen, key=lambda x: int(x, 16)):
    e, s, t, l = seen[a]
    if e >= frac * mx:
        tot += e
        print(f"{a:>8s} {e:11d} {s:6d}  {t:90s} {l[0]}:{l[1]}")
print("sum executed of listed", tot, "max", mx)
