"""Per-source-range instruction and stall-sample totals of a kernel from an .ncu-rep (source page).
usage: ncu_phase_mix.py rep file.cu name:lo-hi [name:lo-hi ...]   (line ranges of `file.cu`)"""
import collections, csv, io, subprocess, sys
rep, fname = sys.argv[1], sys.argv[2]
ranges = [(a.split(':')[0], *map(int, a.split(':')[1].split('-'))) for a in sys.argv[3:]]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
cur_file, hdr = None, None
tot = collections.defaultdict(lambda: [0, 0, 0])      # warp instr, samples, thread instr
ops = collections.defaultdict(lambda: collections.Counter())
line, name = None, None
for r in csv.reader(io.StringIO(src)):
    if not r: continue
    if r[0] == 'File Path': cur_file = r[1]; continue
    if r[0] == 'Line No': hdr = r; continue
    if r[0] == 'Function Name' or hdr is None: continue
    if r[0] != '':
        line = int(r[0]); name = 'other'
        if cur_file.endswith(fname):
            for n, lo, hi in ranges:
                if lo <= line <= hi: name = n
        else:
            name = 'file:' + cur_file.split('/')[-1]
        continue
    try:
        ie, smp, ti = int(r[hdr.index('Instructions Executed')]), int(r[hdr.index('# Samples')]), int(r[hdr.index('Thread Instructions Executed')])
    except ValueError:
        continue
    t = tot[name]; t[0] += ie; t[1] += smp; t[2] += ti
    ops[name][r[3].split()[0].split('.')[0] if r[3].split()[0][0] != '@' else r[3].split()[1].split('.')[0]] += ie
allw = sum(t[0] for t in tot.values()); alls = sum(t[1] for t in tot.values())
for n, t in sorted(tot.items(), key=lambda kv: -kv[1][0]):
    top = ' '.join(f"{k}:{100 * v / t[0]:.0f}%" for k, v in ops[n].most_common(9)) if t[0] else ''
    print(f"{n:18s} warp-instr {t[0]:12d} {100 * t[0] / allw:5.1f}%  samples {100 * t[1] / alls:5.1f}%  | {top}")
print('total warp instructions', allw)
