import csv,sys,collections
fn=sys.argv[1]
rows=list(csv.reader(open(fn)))
cur=None; hdr=None
agg=collections.defaultdict(lambda:[0,0,0,""])
for r in rows:
    if len(r)>=2 and r[0]=="File Path": cur=r[1].split('/')[-1]; continue
    if len(r)>=2 and r[0]=="Function Name": continue
    if len(r)>5 and r[0]=="Line No": hdr=r; ix={}; 
    if len(r)>5 and r[0]=="Line No":
        for k,n in enumerate(hdr):
            if n not in ix: ix[n]=k
        continue
    if hdr is None or len(r)<len(hdr): continue
    try:
        ln=int(r[0]); s=int(r[ix['# Samples']]); i=int(r[ix['Instructions Executed']]); t=int(r[ix['Thread Instructions Executed']])
    except ValueError: continue
    a=agg[(cur,ln)]; a[0]+=s; a[1]+=i; a[2]+=t; a[3]=r[1].strip()[:100]
tot=[sum(a[k] for a in agg.values()) for k in range(3)]
print("total samples %d inst %d thread-inst %d  avg thr/inst %.2f"%(tot[0],tot[1],tot[2],tot[2]/tot[1]))
byfile=collections.defaultdict(lambda:[0,0,0])
for (f,l),a in agg.items():
    for k in range(3): byfile[f][k]+=a[k]
for f,a in byfile.items(): print("%-22s smp %5.1f%% inst %5.1f%% thr/inst %.1f"%(f,100*a[0]/tot[0],100*a[1]/tot[1],a[2]/max(a[1],1)))
n=int(sys.argv[2]) if len(sys.argv)>2 else 40
for (f,l),a in sorted(agg.items(), key=lambda kv:-kv[1][0])[:n]:
    print("%5.1f%% smp %5.1f%% inst thr/inst %5.1f | %s:%d | %s"%(100*a[0]/tot[0],100*a[1]/tot[1],a[2]/max(a[1],1),f,l,a[3]))
