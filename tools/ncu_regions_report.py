import csv,sys,subprocess
rep=sys.argv[1]
raw=subprocess.run(["ncu","-i",rep,"--page","raw","--csv"],stdout=subprocess.PIPE,text=True).stdout
rows=list(csv.reader(raw.splitlines()))
hdr,units,vals=rows[0],rows[1],rows[2]
keys=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','launch__registers_per_thread','sm__warps_active.avg.pct_of_peak_sustained_active','l1tex__t_sector_hit_rate.pct','lts__t_sector_hit_rate.pct','smsp__inst_executed.sum','sm__throughput.avg.pct_of_peak_sustained_elapsed','smsp__average_warp_latency_per_inst_issued.ratio','smsp__thread_inst_executed_per_inst_executed.ratio','sm__cycles_active.avg','smsp__issue_active.avg.pct_of_peak_sustained_active','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','launch__occupancy_limit_shared_mem','launch__occupancy_limit_registers','sm__inst_executed_pipe_alu.sum','smsp__inst_executed_pipe_alu.sum']
for i,h in enumerate(hdr):
    if h in keys: print(h, units[i], vals[i])
src=subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","cuda,sass"],stdout=subprocess.PIPE,text=True).stdout
rows=list(csv.reader(src.splitlines()))
cur=None; hdr=None; out=[]
for r in rows:
    if len(r)==2 and r[0]=="File Path": cur=r[1].split('/')[-1]; continue
    if len(r)>3 and r[0]=="Line No": hdr=r; continue
    if hdr and len(r)==len(hdr) and r[0]!="":
        d=dict(zip(hdr,r))
        try: out.append((cur,int(r[0]),r[1][:100],int(d["# Samples"]),int(d["Instructions Executed"]),d))
        except: pass
tot_s=sum(o[3] for o in out); tot_i=sum(o[4] for o in out)
print("total samples",tot_s,"inst",tot_i)
import re
regions=eval(open(sys.argv[2]).read()) if len(sys.argv)>2 else []
def region(f,l):
    for (ff,a,b,name) in regions:
        if f==ff and a<=l<=b: return name
    return f
reg={}
for o in out:
    k=region(o[0],o[1]); a=reg.setdefault(k,[0,0]); a[0]+=o[3]; a[1]+=o[4]
for k,(s,i) in sorted(reg.items(), key=lambda x:-x[1][0]):
    print("%-40s %5.1f%% samples %5.1f%% inst"%(k,100*s/tot_s,100*i/tot_i))
out.sort(key=lambda o:-o[3])
stalls=[k for k in hdr if k.startswith("stall_") and "Not Issued" not in k]
for o in out[:int(sys.argv[3]) if len(sys.argv)>3 else 30]:
    d=o[5]
    top=sorted(((int(d[k]),k) for k in stalls),reverse=True)[:2]
    print("%-14s %4d %5.1f%% smp %5.1f%% inst %-22s %s"%(o[0][:14],o[1],100*o[3]/tot_s,100*o[4]/tot_i,",".join("%s:%d%%"%(k[6:],100*v/max(o[3],1)) for v,k in top),o[2][:80]))
