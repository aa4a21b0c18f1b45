import csv,re,collections,sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[1]
data=[]
for r in rows[2:]:
    if len(r)!=len(hdr) or not r[0].startswith('0x'):
        if data: break
        continue
    data.append(r)
iS=hdr.index("Source"); iE=hdr.index("Instructions Executed"); iSm=hdr.index("Warp Stall Sampling (All Samples)")
tot=sum(int(r[iE]) for r in data); totS=sum(int(r[iSm]) for r in data)
print("total instr",tot,"samples",totS,"lines",len(data))
h=collections.Counter(); hs=collections.Counter()
for r in data:
    t=r[iS].strip().split()
    op=t[1] if t[0].startswith('@') else t[0]
    op=op.split('.')[0]
    h[op]+=int(r[iE]); hs[op]+=int(r[iSm])
for op,c in h.most_common(32): print(f"{op:10s} {c:9d} {c/tot*100:5.1f}%  samples {hs[op]/totS*100:5.1f}%")
if len(sys.argv)>2:
    d=sorted(enumerate(data),key=lambda x:-int(x[1][iSm]))[:int(sys.argv[2])]
    for i,r in d: print(i,r[iSm],r[iE],r[iS][:90])
