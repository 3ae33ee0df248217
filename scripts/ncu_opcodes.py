"""Executed-instruction and stall-sample shares per SASS opcode from an `ncu --page source --csv` export.
usage: ncu_opcodes.py source.csv [top]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, body = rows[1], rows[2:]
iE, iS, iN = hdr.index('Instructions Executed'), hdr.index('Source'), hdr.index('# Samples')
cnt, smp = collections.Counter(), collections.Counter()
for r in body:
    parts = r[iS].strip().split()
    op = parts[1] if parts[0].startswith('@') else parts[0]
    op = '.'.join(op.split('.')[:3]) if op.startswith(('LDS', 'LDG', 'STG', 'STS', 'F2F', 'LDGSTS')) else op.split('.')[0]
    cnt[op] += int(r[iE])
    smp[op] += int(r[iN])
tot, ts = sum(cnt.values()), sum(smp.values())
print(rows[0][1])
for op, c in cnt.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 30):
    print(f'{op:22s} {c / tot * 100:6.2f}% of warp instructions  {smp[op] / ts * 100:6.2f}% of stall samples')
