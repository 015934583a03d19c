import csv,sys,collections
rows=list(csv.reader(open(sys.argv[1])))
cur=None; hdr=None
agg=collections.OrderedDict()
for r in rows:
    if len(r)>=2 and r[0]=='File Path': cur=r[1].split('/')[-1]; continue
    if len(r)>=2 and r[0]=='Line No': hdr=r; ie=hdr.index('Instructions Executed'); isamp=hdr.index('# Samples'); continue
    if hdr and len(r)==len(hdr):
        if not r[0].strip().isdigit(): continue
        try: n=int(r[ie]); s=int(r[isamp])
        except: continue
        key=(cur,int(r[0]))
        if key not in agg: agg[key]=[0,0,r[1].strip()[:95],0]
        agg[key][0]+=n; agg[key][1]+=s; agg[key][3]+=1
tot=sum(v[0] for v in agg.values()); ts=sum(v[1] for v in agg.values())
print('total inst',tot,'samples',ts)
pf=collections.Counter()
for (f,l),v in agg.items(): pf[f]+=v[0]
print([(f,round(100*n/tot,1)) for f,n in pf.most_common()])
top=int(sys.argv[2]) if len(sys.argv)>2 else 50
for (f,l),v in sorted(agg.items(), key=lambda kv:-kv[1][0])[:top]:
    print(f"{100*v[0]/tot:5.2f}% samp {100*v[1]/max(ts,1):5.2f}% sass {v[3]:4d} {f}:{l}  {v[2]}")
