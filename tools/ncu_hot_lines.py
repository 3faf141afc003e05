import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
src = subprocess.run(["ncu","-i",rep,"-k","regex:"+kern,"--page","source","--csv","--print-source","cuda,sass"],capture_output=True,text=True).stdout
rows=list(csv.reader(src.splitlines()))
hi=next(i for i,r in enumerate(rows) if r and r[0]=="Line No"); hdr=rows[hi]
lines=[r for r in rows[hi+1:] if len(r)>8 and r[2]=="-" and r[4].isdigit()]
tot=sum(int(r[4]) for r in lines); print("total samples",tot)
sidx={n:hdr.index(n) for n in hdr if n.startswith("stall_") and "Not Issued" not in n}
lines.sort(key=lambda r:-int(r[4]))
for r in lines[:int(sys.argv[3]) if len(sys.argv)>3 else 22]:
    st={k[6:]:int(r[i]) for k,i in sidx.items() if i<len(r) and r[i].isdigit() and int(r[i])>0}
    t3=sorted(st.items(),key=lambda x:-x[1])[:3]
    print("%5.1f%% L%-4s %-90s %s"%(100*int(r[4])/tot, r[0], r[1].strip()[:90], t3))
raw=subprocess.run(["ncu","-i",rep,"-k","regex:"+kern,"--page","raw","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines())); h,u,v=rows[0],rows[1],rows[2]
for name in ["gpu__time_duration.sum","dram__bytes_read.sum","dram__bytes_write.sum","sm__warps_active.avg.pct_of_peak_sustained_active","smsp__issue_active.avg.pct_of_peak_sustained_active","launch__registers_per_thread","launch__occupancy_limit_shared_mem","launch__cluster_max_active" ,"sm__throughput.avg.pct_of_peak_sustained_elapsed","lts__throughput.avg.pct_of_peak_sustained_elapsed","l1tex__throughput.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active","smsp__inst_executed.sum","sm__cycles_active.avg"]:
    if name in h: i=h.index(name); print("%-70s %-10s %s"%(name,u[i],v[i]))
