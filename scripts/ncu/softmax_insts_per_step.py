import csv, sys, collections
rows=list(csv.reader(open(sys.argv[1]))); hdr=rows[1]; data=rows[2:]
col={n:i for i,n in enumerate(hdr)}
num=lambda x: float(x.replace(',','')) if x.replace(',','').replace('.','',1).isdigit() else 0.0
idx=[i for i,r in enumerate(data) if 'USETMAXREG' in r[col['Source']]]
a=idx[-1]
steps=sum(num(r[col['Instructions Executed']]) for r in data[a:] if 'LDTM.x32' in r[col['Source']])/4
c=collections.Counter()
for r in data[a:]:
    t=r[col['Source']].strip().split()
    op=t[1] if t and t[0].startswith('@') else (t[0] if t else '?')
    c[op]+=num(r[col['Instructions Executed']])
print('warp-steps',steps)
for k,v in c.most_common(45): print(f'{k:28s} {v/steps:7.1f} per step')
