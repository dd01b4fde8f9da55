import csv,sys,subprocess
rep=sys.argv[1]; fname=sys.argv[2]; a=int(sys.argv[3]); b=int(sys.argv[4])
src=subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","cuda,sass"],stdout=subprocess.PIPE,text=True).stdout
rows=list(csv.reader(src.splitlines()))
cur=None; hdr=None; out=[]
for r in rows:
    if len(r)==2 and r[0]=="File Path": cur=r[1].split('/')[-1]; continue
    if len(r)>3 and r[0]=="Line No": hdr=r; continue
    if hdr and len(r)==len(hdr) and r[0]!="":
        d=dict(zip(hdr,r))
        try: out.append((cur,int(r[0]),r[1][:110],int(d["# Samples"]),int(d["Instructions Executed"])))
        except: pass
tot_s=sum(o[3] for o in out); tot_i=sum(o[4] for o in out)
n=float(sys.argv[5]) if len(sys.argv)>5 else 1.0
for o in out:
    if o[0]==fname and a<=o[1]<=b and o[4]>0:
        print("%4d %5.1f%%s %6.1f i/u  %s"%(o[1],100*o[3]/tot_s,o[4]/n,o[2].strip()[:100]))
