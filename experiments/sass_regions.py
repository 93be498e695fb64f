"""Summarise an `ncu --page source --csv` export of conv1d_adain_tc_kernel: split the SASS into regions at
branch targets that look like role boundaries and print, per contiguous hot region, samples / instructions
executed / dominant stall reasons."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; data = [r for r in rows[2:] if len(r) == len(hdr) and r[0].startswith("0x")]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
W = int(sys.argv[2]) if len(sys.argv) > 2 else 64      # window of instructions
tot = sum(int(r[ix['# Samples']]) for r in data)
totex = sum(int(r[ix['Instructions Executed']]) for r in data)
print("total samples", tot, "instr executed", totex, "n sass", len(data))
for s in range(0, len(data), W):
    blk = data[s:s + W]
    smp = sum(int(r[ix['# Samples']]) for r in blk)
    ex = sum(int(r[ix['Instructions Executed']]) for r in blk)
    if smp < tot * 0.01 and ex < totex * 0.01:
        continue
    st = collections.Counter()
    for r in blk:
        for h in stalls:
            st[h[6:]] += int(r[ix[h]] or 0)
    ops = collections.Counter(r[ix['Source']].split()[0 if not r[ix['Source']].strip().startswith('@') else 1].split('.')[0] for r in blk)
    tag = ','.join(k for k in ('MUFU', 'UTCHMMA', 'LDTM', 'UTMALDG', 'UTMASTG', 'SYNCS', 'STS', 'LDS', 'STG', 'DADD', 'RED', 'ATOM') if ops.get(k))
    print(f"[{s:5d}] smp {smp / tot * 100:5.1f}%  exec {ex / totex * 100:5.1f}%  {tag:40s} " + ' '.join(f"{k}={v / max(smp, 1) * 100:.0f}%" for k, v in st.most_common(4)))
