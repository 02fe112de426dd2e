import csv, sys, subprocess
rep=sys.argv[1]; which=int(sys.argv[2]) if len(sys.argv)>2 else 0
out=subprocess.run(['ncu','-i',rep,'--page','source','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
blocks=[]; cur=None
for r in rows:
    if r and r[0]=='Kernel Name':
        cur={'name':r[1],'rows':[]}; blocks.append(cur)
    elif cur is not None: cur['rows'].append(r)
names=[b['name'][:80] for b in blocks]
print(names)
b=blocks[which]
hdr=b['rows'][0]; data=[r for r in b['rows'][1:] if len(r)==len(hdr)]
isrc=hdr.index('Source'); isamp=hdr.index('# Samples'); iex=hdr.index('Instructions Executed')
stall_cols=[i for i,h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot=sum(int(r[isamp] or 0) for r in data)
print('total samples', tot)
agg={}
for r in data:
    for i in stall_cols: agg[hdr[i]]=agg.get(hdr[i],0)+int(r[i] or 0)
print(sorted(agg.items(), key=lambda kv:-kv[1])[:8])
top=sorted(range(len(data)), key=lambda i:-int(data[i][isamp] or 0))[:int(sys.argv[3]) if len(sys.argv)>3 else 30]
for i in sorted(top):
    r=data[i]
    st=sorted(((int(r[k] or 0),hdr[k][6:]) for k in stall_cols),reverse=True)[:2]
    print(i, r[isamp], r[iex], r[isrc][:64], st)
