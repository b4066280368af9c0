import csv, collections, sys
rows=list(csv.reader(open(sys.argv[1])))
agg=collections.defaultdict(lambda:[0,0,0]); cur=None; hdr=None; inst=0
for r in rows:
    if not r: continue
    if r[0]=="File Path": cur=r[1].split('/')[-1]; hdr=None; continue
    if r[0]=="Kernel Name": inst+=1; continue
    if r[0]=="Function Name": continue
    if r[0]=="Line No": hdr=r; isamp=hdr.index('# Samples'); iex=hdr.index('Instructions Executed'); ith=hdr.index('Thread Instructions Executed'); continue
    if hdr is None or cur is None or r[0]=='' : continue
    try: ln=int(r[0]); s=int(r[isamp] or 0); e=int(r[iex] or 0); t=int(r[ith] or 0)
    except: continue
    a=agg[(cur,ln)]; a[0]+=s; a[1]+=e; a[2]+=t
tot=sum(v[0] for v in agg.values()) or 1; tote=sum(v[1] for v in agg.values()) or 1
print('total samples',tot,'instr',tote)
byfile=collections.defaultdict(lambda:[0,0,0])
for (f,l),v in agg.items():
    b=byfile[f]; b[0]+=v[0]; b[1]+=v[1]; b[2]+=v[2]
for f,v in sorted(byfile.items(), key=lambda x:-x[1][0])[:9]: print(f"{f:24s} {v[0]/tot*100:5.1f}% samples {v[1]/tote*100:5.1f}% instr  avg threads {v[2]/max(v[1],1):.1f}")
print('top lines')
for (f,l),v in sorted(agg.items(), key=lambda x:-x[1][0])[:int(sys.argv[2]) if len(sys.argv)>2 else 22]: print(f"{f}:{l}".ljust(30),f"{v[0]/tot*100:5.1f}% smp", f"{v[1]/tote*100:5.1f}% instr", f"thr {v[2]/max(v[1],1):.1f}")
