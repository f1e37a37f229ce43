import csv,sys,re
from collections import defaultdict
f=sys.argv[1]
rows=list(csv.reader(open(f)))
hdr=rows[1]; data=rows[2:]
ix={h:i for i,h in enumerate(hdr)}
def num(x):
    try: return float(x.replace(',',''))
    except: return 0.0
agg=defaultdict(lambda: defaultdict(float))
stalls=[h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
for r in data:
    if len(r)<len(hdr): continue
    op=r[ix['Source']].strip()
    op=re.sub(r'^@!?U?P\d+\s+','',op).split()[0]
    base=op.split('.')[0]
    if base in('LDS','STS','LDG','STG'):
        base=op if op.count('.')<2 else '.'.join(op.split('.')[:2])
    a=agg[base]
    a['inst']+=num(r[ix['Instructions Executed']]); a['samples']+=num(r[ix['# Samples']])
    for s in stalls: a[s]+=num(r[ix[s]])
tot_i=sum(a['inst'] for a in agg.values()); tot_s=sum(a['samples'] for a in agg.values())
print(f"total inst {tot_i:.3e} samples {tot_s:.0f}")
print(f"{'op':14s} {'inst%':>6s} {'smp%':>6s}  top stalls (share of all samples)")
for op,a in sorted(agg.items(), key=lambda kv:-kv[1]['samples'])[:18]:
    top=sorted(((a[s],s) for s in stalls), reverse=True)[:4]
    print(f"{op:14s} {100*a['inst']/tot_i:6.2f} {100*a['samples']/tot_s:6.2f}  "+", ".join(f"{s[6:]} {100*v/tot_s:.1f}" for v,s in top if v>0))
