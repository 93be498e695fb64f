#!/usr/bin/env python
"""Per-group table of the fused conv launches of one forward from an `ncu --csv --metrics ...` log
(experiments/ncu_step.sh): python experiments/ncu_groups.py gpurun_out/r02_fused_metrics.csv profiles/r02_fused_groups.txt"""
import collections, csv, json, os, sys
src, dst = sys.argv[1], sys.argv[2]
rows = [r for r in csv.reader(open(src)) if len(r) > 10]
hdr = rows[0]; ix = {h: i for i, h in enumerate(hdr)}
by = collections.OrderedDict()
for r in rows[1:]:
    by.setdefault(r[ix['ID']], {})[r[ix['Metric Name']]] = float(r[ix['Metric Value']].replace(',', ''))
    by[r[ix['ID']]]['name'] = r[ix['Kernel Name']]
L = list(by.values())
assert len(L) == 49, len(L)
names = ['s0.noise_res k7'] * 6 + ['s0.rb0 k3'] * 6 + ['s0.rb1 k7'] * 6 + ['s0.rb2 k11'] * 6 + ['s1.noise_res k11'] * 6 + \
        ['s1.rb3 k3'] * 6 + ['s1.rb4 k7'] * 6 + ['s1.rb5 k11'] * 6 + ['conv_post k7']
g = collections.OrderedDict(); tot = tw = 0
for n, m in enumerate(L):
    key = names[n] + (' conv1' if n % 2 == 0 else ' conv2') if n < 48 else names[n]
    t = m['gpu__time_duration.sum'] / 1e3; tp = m['sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed']
    form = 'T' if ', 1, 1>' in m['name'].split('(')[0] or ', true' in m['name'] else 'R'
    a = g.setdefault(key, [0, 0, 0, 0, 0, set()]); a[0] += 1; a[1] += t; a[2] += t * tp
    a[3] += m['dram__bytes_read.sum']; a[4] += m['dram__bytes_write.sum']; a[5].add(m['name'].split('<')[1].split('>')[0])
    tot += t; tw += t * tp
out = ["Round 2, final build - the 49 fused launches (conv1d_adain_tc_kernel) of one forward, B=64 x T=320, ncu --clock-control none",
       "(experiments/ncu_step.sh; template arguments <MT, BN, CB, SWAP, CL>)",
       "group                      n   us/launch  tensor-pipe %   DRAM read GB  write GB   TB/s   instance"]
for k, a in g.items():
    out.append(f"{k:26s} {a[0]}   {a[1]/a[0]:8.1f}   {a[2]/a[1]:6.1f}        {a[3]/a[0]/1e9:6.3f}      {a[4]/a[0]/1e9:6.3f}   {(a[3]+a[4])/a[1]/1e6:5.2f}   {' / '.join(sorted(a[5]))}")
out.append(f"total {tot/1e3:.2f} ms; time-weighted sm__pipe_tensor_cycles_active = {tw/tot:.1f} %")
open(dst, 'w').write('\n'.join(out) + '\n'); print('\n'.join(out))
json.dump({"traffic_bytes_per_launch": sum(m['dram__bytes_read.sum'] + m['dram__bytes_write.sum'] for m in L) / len(L), "launches": len(L),
           "source": "profiles/r02_fused_metrics.csv"}, open(os.path.join(os.path.dirname(dst), 'r02_fused_traffic.json'), 'w'))
