"""Stall-sample share per SASS opcode and the hottest SASS instructions of an .ncu-rep (de-duplicated by address).
usage: ncu_sass_hot.py rep [n]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
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
    try: seen[r[2]] = (int(r[6]), int(r[7]), r[3].strip()[:64], line)
    except ValueError: continue
tot = sum(v[0] for v in seen.values()); ex = sum(v[1] for v in seen.values())
byop, exop = collections.Counter(), collections.Counter()
for s, e, t, l in seen.values():
    w = t.split(); o = (w[1] if w[0].startswith('@') else w[0]).split('.')[0]
    byop[o] += s; exop[o] += e
print(f"samples {tot}, warp instructions {ex}")
for k, v in byop.most_common(22): print(f"{k:14s} samples {100 * v / tot:5.1f}%   executed {100 * exop[k] / ex:5.1f}%")
for s, e, t, l in sorted(seen.values(), reverse=True)[:n]: print(f"{100 * s / tot:5.1f}%  exec {e:10d}  {t:64s} {l[0]}:{l[1]}")
