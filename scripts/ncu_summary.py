import csv, sys, subprocess
rep=sys.argv[1]
raw=subprocess.run(["ncu","-i",rep,"--page","raw","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines()))
hdr,units,vals=rows[0],rows[1],rows[2]
want=['gpu__time_duration.sum','dram__bytes_read.sum ','dram__bytes_write.sum ','launch__registers_per_thread','launch__occupancy_limit','launch__waves','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum ','sm__inst_executed.avg.per_cycle_elapsed','smsp__issue_active.avg.pct','sass__inst_executed_local','smsp__thread_inst_executed_per_inst_executed.ratio','smsp__average_warps_issue_stalled','sm__inst_executed_pipe_fma.avg.pct','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum ','l1tex__t_sector_pipe_lsu_mem_local_op_ld_hit_rate','smsp__inst_executed_op_shared','l1tex__data_bank_conflicts_pipe_lsu_mem_shared','launch__shared_mem_per_block','launch__block_size','launch__grid_size','sm__throughput.avg.pct']
for h,u,v in zip(hdr,units,vals):
    if any((w in h+' ') for w in want):
        try:
            if 'stalled' in h and float(v)<0.3: continue
        except: pass
        print(h,u,v)
src=subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","cuda,sass"],capture_output=True,text=True).stdout
rows=list(csv.reader(src.splitlines()))
cur=None; agg={}; hdr=None
for r in rows:
    if len(r)>=2 and r[0]=="File Path": cur=r[1].split('/')[-1]; continue
    if len(r)>3 and r[0]=="Line No": hdr=r; continue
    if hdr is None or len(r)<10: continue
    if r[0]!="":
        try: ln=int(r[0]); s=int(r[6]); ie=int(r[7]); te=int(r[8])
        except: continue
        agg[(cur,ln)]=(s,ie,te,r[1].strip()[:90])
tot_s=sum(v[0] for v in agg.values()); tot_i=sum(v[1] for v in agg.values())
print("samples",tot_s,"inst",tot_i)
byfile={}
for (f,ln),(s,ie,te,_) in agg.items():
    d=byfile.setdefault(f,[0,0,0]); d[0]+=s; d[1]+=ie; d[2]+=te
for f,(s,ie,te) in sorted(byfile.items(), key=lambda kv:-kv[1][0]): print(f"{f:22s} samples {100*s/tot_s:5.1f}% inst {100*ie/tot_i:5.1f}% avgthr {te/max(ie,1):.1f}")
N=int(sys.argv[2]) if len(sys.argv)>2 else 30
for (f,ln),(s,ie,te,srcl) in sorted(agg.items(), key=lambda kv:-kv[1][0])[:N]:
    print(f"{f:18s}:{ln:4d} s {100*s/tot_s:5.1f}% i {100*ie/tot_i:5.1f}% thr {te/max(ie,1):5.1f} | {srcl}")
