import csv, sys, re, collections
path=sys.argv[1]
rows=list(csv.reader(open(path)))
hdr=rows[1]; data=rows[2:]
col={n:i for i,n in enumerate(hdr)}
stall=[n for n in hdr if n.startswith('stall_') and 'Not Issued' not in n]
def num(x):
    try: return float(x.replace(',',''))
    except: return 0.0
# locate role regions by USETMAXREG instructions
idx=[i for i,r in enumerate(data) if 'USETMAXREG' in r[col['Source']]]
print('USETMAXREG at', idx, [data[i][col['Source']].strip()[:60] for i in idx])
tot=sum(num(r[col['# Samples']]) for r in data)
print('total samples', tot)
def region(a,b,name):
    agg=collections.Counter(); n=0; inst=0
    for r in data[a:b]:
        s=num(r[col['# Samples']]); n+=s; inst+=num(r[col['Instructions Executed']])
        for k in stall: agg[k]+=num(r[col[k]])
    print(f'--- {name}: rows {a}-{b}, samples {n:.0f} ({100*n/tot:.1f} %), warp-instructions {inst:.0f}')
    for k,v in agg.most_common(8): print(f'    {k:28s} {v:8.0f} {100*v/max(n,1):5.1f} %')
    # top instructions
    top=sorted(data[a:b], key=lambda r:-num(r[col['# Samples']]))[:int(sys.argv[2]) if len(sys.argv)>2 else 14]
    for r in top:
        ss={k:num(r[col[k]]) for k in stall}
        k=max(ss,key=ss.get)
        print(f"    {num(r[col['# Samples']]):7.0f}  {r[col['Source']].strip()[:70]:70s} {k} {ss[k]:.0f}")
bounds=idx+[len(data)]
region(0,idx[0],'prologue')
for a,b in zip(bounds[:-1],bounds[1:]):
    region(a,b,data[a][col['Source']].strip()[:40])
