"""Attribute an ncu SASS source page to CUDA source lines.
usage: ncu_lines.py sass.csv disasm.txt kernel_substr source.cu"""
import csv, re, collections, sys
sass_csv, dis, ksub, srcf = sys.argv[1:5]
cur = None; seq = []; active = False
for ln in open(dis):
    if ln.startswith('//--------------------- .text'):
        active = ksub in ln; continue
    if ln.startswith('//--------------------- ') and '.text' not in ln:
        active = False
    if not active: continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+\S', ln):
        seq.append((cur, ln.strip()))
rows = list(csv.reader(open(sass_csv)))
hdr = rows[1]; body = rows[2:]
print(len(seq), len(body))
iE = hdr.index('Instructions Executed'); iN = hdr.index('# Samples')
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(int(r[iE]) for r in body); totS = sum(int(r[iN]) for r in body)
agg = collections.Counter(); aggS = collections.Counter(); st = collections.defaultdict(collections.Counter)
for (loc, txt), r in zip(seq, body):
    agg[loc] += int(r[iE]); aggS[loc] += int(r[iN])
    for i in stall_cols:
        v = int(r[i] or 0)
        if v: st[loc][hdr[i][6:]] += v
import os
srcdir = os.path.dirname(srcf)
srcs = {f: open(os.path.join(srcdir, f)).read().split('\n') for f in os.listdir(srcdir) if f.endswith(('.cu', '.cuh', '.inl'))}
tots = collections.Counter()
for loc in st:
    for k, v in st[loc].items(): tots[k] += v
print('stall totals:', ', '.join(f'{k} {v/totS*100:.1f}%' for k, v in tots.most_common(10)))
for key, c in sorted(aggS.items(), key=lambda kv: -kv[1])[:int(sys.argv[5]) if len(sys.argv) > 5 else 40]:
    f, l = key if key else ('?', 0)
    s = srcs[f][l - 1].strip()[:90] if f in srcs and 0 < l <= len(srcs[f]) else f
    top = ', '.join(f'{k}:{v/totS*100:.1f}' for k, v in st[key].most_common(3))
    print(f'{f[:12]:12s}:{l:4d} inst {agg[key]/tot*100:5.2f}% smp {c/totS*100:5.2f}% [{top}]  {s}')
