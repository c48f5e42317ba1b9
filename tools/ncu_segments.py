import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[1]; data=rows[2:]
isrc=hdr.index('Source'); iex=hdr.index('Instructions Executed'); ismp=hdr.index('# Samples')
stall_cols=[i for i,h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot=sum(int(r[iex]) for r in data); ts=sum(int(r[ismp]) for r in data); print('total',tot, 'samples', ts)
prev=None; start=0; segs=[]
for i,r in enumerate(data):
    n=int(r[iex]); lvl = 0 if n==0 else len(str(n))
    if prev is not None and lvl!=prev: segs.append((start,i)); start=i
    prev=lvl
segs.append((start,len(data)))
for a,b in segs:
    t=sum(int(r[iex]) for r in data[a:b]); sm=sum(int(r[ismp]) for r in data[a:b])
    if t>5e6 or sm>1500:
        st={hdr[c]:sum(int(r[c] or 0) for r in data[a:b]) for c in stall_cols}
        top=sorted(st.items(), key=lambda kv:-kv[1])[:3]
        print(a,b,'instr',t,'samples',sm, f'{100*sm/ts:.1f}%', data[a][isrc][:40], '| n=', data[a][iex], top)
best=sorted(data,key=lambda r:-int(r[ismp]))[:10]
for r in best: print(r[ismp], r[iex], r[isrc][:90])
