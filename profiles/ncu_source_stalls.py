import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
which=int(sys.argv[2]) if len(sys.argv)>2 else 0
topn=int(sys.argv[3]) if len(sys.argv)>3 else 25
blocks=[];cur=None
for r in rows:
    if r and r[0]=='Kernel Name': cur=[r]; blocks.append(cur)
    elif cur is not None: cur.append(r)
b=blocks[which]
hdr=b[1]
ia=hdr.index('Source'); isamp=hdr.index('# Samples'); ie=hdr.index('Instructions Executed')
stalls=[h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
body=[r for r in b[2:] if len(r)==len(hdr)]
tot=sum(int(r[isamp]) for r in body)
print(b[0][1][:50],'samples',tot,'instr',sum(int(r[ie]) for r in body))
agg={}
for r in body:
    for s in stalls: agg[s]=agg.get(s,0)+int(r[hdr.index(s)])
print([(k,round(v/tot,3)) for k,v in sorted(agg.items(), key=lambda x:-x[1])[:7]])
for r in sorted(body, key=lambda r:-int(r[isamp]))[:topn]:
    st=sorted(((int(r[hdr.index(s)]),s) for s in stalls), reverse=True)[:2]
    print(f"{int(r[isamp])/tot:6.1%} {int(r[ie])/1e6:7.2f}M {r[ia].strip()[:56]:56s} {st}")
