import csv, subprocess, sys
rep, kern, skip = sys.argv[1], sys.argv[2], sys.argv[3]
out = subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","cuda,sass","--kernel-name","regex:"+kern,"--launch-skip",skip,"--launch-count","1"],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=None; lines={}; cur=None; fname=None
for r in rows:
    if r and r[0]=="File Path": fname=r[1].split('/')[-1]; continue
    if r and r[0]=="Line No": hdr=r; si=hdr.index("Warp Stall Sampling (All Samples)"); continue
    if hdr is None or len(r)<=si: continue
    if r[0]!='':
        try: n=int(r[si])
        except: continue
        lines[(fname,int(r[0]),r[1][:100])]=n
tot=sum(lines.values()); print("total samples",tot)
for k,v in sorted(lines.items(), key=lambda kv:-kv[1])[:int(sys.argv[4]) if len(sys.argv)>4 else 22]:
    print("%5d %5.1f%%  %s:%d  %s"%(v,100*v/tot,k[0],k[1],k[2]))
