import csv, sys, collections
rows=list(csv.reader(open(sys.argv[1]))); hdr=rows[1]; data=rows[2:]
col={n:i for i,n in enumerate(hdr)}
stall=[n for n in hdr if n.startswith('stall_') and 'Not Issued' not in n]
num=lambda x: float(x.replace(',','')) if x.replace(',','').replace('.','',1).isdigit() else 0.0
idx=[i for i,r in enumerate(data) if 'USETMAXREG' in r[col['Source']]]
a=idx[-1]
reg=data[a:]
byop=collections.Counter(); bystall=collections.Counter(); inst=collections.Counter()
tot=0
for r in reg:
    src=r[col['Source']].strip(); s=num(r[col['# Samples']])
    toks=src.split()
    op=toks[1] if toks and toks[0].startswith('@') else (toks[0] if toks else '?')
    op=op.split('.')[0]
    if op in('BRA','EXIT','SYNCS','NANOSLEEP'): op='wait-loop/'+op
    byop[op]+=s; inst[op]+=num(r[col['Instructions Executed']]); tot+=s
    if not op.startswith('wait-loop'):
        for k in stall: bystall[k]+=num(r[col[k]])
print('softmax-region samples',tot)
for k,v in byop.most_common(25): print(f'  {k:22s} samples {v:6.0f} {100*v/tot:5.1f} %   warp-insts {inst[k]:10.0f}')
w=sum(bystall.values())
print('work (non-wait-loop) samples',w)
for k,v in bystall.most_common(10): print(f'  {k:26s} {v:6.0f} {100*v/w:5.1f} %')
print('total warp-instructions in region', sum(inst.values()))
