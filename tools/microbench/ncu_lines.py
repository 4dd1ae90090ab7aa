import csv,sys,subprocess
rep,kern=sys.argv[1],sys.argv[2]
top=int(sys.argv[3]) if len(sys.argv)>3 else 40
out=subprocess.run(["ncu","-i",rep,"--page","source","--csv","--kernel-name","regex:"+kern,"--print-source","cuda,sass"],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
h=None; cur=None; agg=[]; tot=0; ts=0
for r in rows:
    if len(r)>=2 and r[0] in('File Path','File Name'): cur=r[1].split('/')[-1]; continue
    if 'Instructions Executed' in r and 'Address' in r:
        h=r; iE=h.index('Instructions Executed'); iSm=h.index('# Samples'); iA=h.index('Address'); continue
    if h is None or len(r)<len(h): continue
    if r[iA]!='-' : continue
    try: e=int(r[iE]); sm=int(r[iSm])
    except: continue
    agg.append((e,sm,cur,r[0],r[1].strip())); tot+=e; ts+=sm
print('total',tot,'samples',ts)
for e,sm,f,l,src in sorted(agg,key=lambda x:-x[0])[:top]:
    print(f"{e:11d} {100*e/tot:5.1f}% smp {100*sm/max(ts,1):5.1f}%  {f}:{l:>4s}  {src[:100]}")
