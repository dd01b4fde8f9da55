import csv,sys,subprocess,collections
rep=sys.argv[1]
src=subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","cuda,sass"],stdout=subprocess.PIPE,text=True).stdout
rows=list(csv.reader(src.splitlines()))
hdr=None; tot=collections.Counter(); n=0
for r in rows:
    if len(r)>3 and r[0]=="Line No": hdr=r; continue
    if hdr and len(r)==len(hdr) and r[0]!="":
        d=dict(zip(hdr,r))
        try:
            n+=int(d["# Samples"])
            for k in hdr:
                if k.startswith("stall_") and "Not Issued" not in k: tot[k]+=int(d[k])
        except: pass
for k,v in tot.most_common(12): print("%-30s %5.1f%%"%(k,100*v/n))
