import csv, subprocess, sys
rep, kern, skip, lo, hi = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5])
out = subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","cuda,sass","--kernel-name","regex:"+kern,"--launch-skip",skip,"--launch-count","1"],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=None; cur=None; fname=None
for r in rows:
    if r and r[0]=="File Path": fname=r[1].split('/')[-1]; continue
    if r and r[0]=="Line No": hdr=r; si=hdr.index("Warp Stall Sampling (All Samples)"); b=hdr.index("stall_barrier"); e=hdr.index("stall_wait"); ie=hdr.index("Instructions Executed"); continue
    if hdr is None or len(r)<=si: continue
    if r[0]!='':
        cur=int(r[0]) if r[0].isdigit() else None; continue
    if fname!=sys.argv[6] or cur is None or not (lo<=cur<=hi): continue
    try: n=int(r[si])
    except: n=0
    reasons={hdr[j][6:]:int(r[j]) for j in range(b,e+1) if r[j] not in ('','0','-')}
    print("%4d %5d x%-7s %-70s %s"%(cur,n,r[ie],r[3][:70],reasons if n else ''))
