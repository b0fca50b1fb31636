import csv,collections,sys
fn=sys.argv[1]; srcfile=sys.argv[2]
rows=list(csv.reader(open(fn)))
cur=None;hdr=None;curln=None
amap={}
for r in rows:
    if len(r)>=2 and r[0]=="File Path": cur=r[1].split('/')[-1]; continue
    if len(r)>5 and r[0]=="Line No":
        hdr=r; ix={}
        for k,n in enumerate(hdr):
            if n not in ix: ix[n]=k
        continue
    if hdr is None or len(r)<len(hdr): continue
    if r[0]!="":
        try: curln=int(r[0])
        except ValueError: pass
        continue
    addr=r[2]
    if addr.startswith("0x"):
        try: amap[int(addr,16)]=(cur,curln,int(r[ix['# Samples']]),int(r[ix['Instructions Executed']]),int(r[ix['Thread Instructions Executed']]), int(r[ix['stall_no_inst']]))
        except ValueError: pass
src=open(srcfile).read().splitlines()
names=sys.argv[3:]
marks=[]
for i,l in enumerate(src,1):
    for name in names:
        if name in l: marks.append((i,name))
base=srcfile.split('/')[-1]
def region(ln):
    r="prologue"
    for i,name in marks:
        if ln>=i: r=name
    return r
stat=collections.OrderedDict()
curreg="prologue"
for a in sorted(amap):
    f,ln,smp,ins,thr,noi=amap[a]
    if f==base: curreg=region(ln)
    s=stat.setdefault(curreg,[0,0,0,0,0]); s[0]+=1; s[1]+=smp; s[2]+=ins; s[3]+=thr; s[4]+=noi
ts=sum(s[1] for s in stat.values()); ti=sum(s[2] for s in stat.values())
for k,s in stat.items(): print("%-22s sass %5d (%4.0f KB)  smp %5.1f%%  warp-inst %5.1f%%  thr/inst %5.1f  no_inst %4.0f%% of its samples"%(k,s[0],s[0]*16/1024,100*s[1]/ts,100*s[2]/ti,s[3]/max(s[2],1),100*s[4]/max(s[1],1)))
