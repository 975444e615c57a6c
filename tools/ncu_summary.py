"""Summarise one ncu --set full capture: python tools/ncu_summary.py raw.csv source.csv
(raw.csv = ncu -i X.ncu-rep --page raw --csv, source.csv = --page source --csv).  Prints the unit
utilisations, the stall mix from the warp samples, per-opcode sample shares, the L1TEX cost of the
memory instructions and the hottest instructions."""
import csv,sys
raw,srcf=sys.argv[1],sys.argv[2]
rows=list(csv.reader(open(raw)))
hdr=rows[0]; units=rows[1]; vals=rows[2]
keys=["gpu__time_duration.sum","dram__bytes_read.sum","dram__bytes_write.sum","launch__registers_per_thread","launch__occupancy_limit_registers","launch__occupancy_limit_shared_mem","sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active","smsp__issue_active.avg.pct_of_peak_sustained_active","smsp__inst_executed.sum","l1tex__data_pipe_lsu_wavefronts_mem_shared.sum","l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum","lts__t_sector_hit_rate.pct","sm__cycles_active.avg","lts__throughput.avg.pct_of_peak_sustained_elapsed","l1tex__throughput.avg.pct_of_peak_sustained_elapsed","l1tex__m_xbar2l1tex_read_bytes.sum","l1tex__data_pipe_lsu_wavefronts.sum","l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed","l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed","l1tex__t_requests_pipe_lsu.sum","l1tex__f_wavefronts.sum"]
for i,h in enumerate(hdr):
    if h in keys: print(h, units[i], vals[i])
rows=list(csv.reader(open(srcf)))
hdr=rows[1]; data=[r for r in rows[2:] if len(r)==len(hdr) and r[0]!="Address"]
ix={h:i for i,h in enumerate(hdr)}
tot=sum(int(r[ix["# Samples"]] or 0) for r in data)
from collections import defaultdict
byop=defaultdict(lambda:[0,0]); stall=defaultdict(int); agg=defaultdict(lambda: defaultdict(float))
for r in data:
    src=r[ix["Source"]].split()
    op=(src[1] if src[0].startswith("@") else src[0]).split('.')[0]
    n=int(r[ix["# Samples"]] or 0); byop[op][0]+=n; byop[op][1]+=int(r[ix["Instructions Executed"]] or 0)
    for k in ["stall_lg","stall_long_sb","stall_math","stall_mio","stall_short_sb","stall_wait","stall_not_selected","stall_selected","stall_dispatch","stall_branch_resolving","stall_no_inst"]:
        stall[k]+=int(r[ix[k]] or 0)
    if op in ("STG","LDGSTS","LDS","LDG","UTMALDG","SHFL"):
        for k in ["Instructions Executed","L1 Tag Requests Global","L1 Wavefronts Shared","L1 Wavefronts Shared Ideal","L2 Theoretical Sectors Global","L2 Theoretical Sectors Global Ideal"]:
            try: agg[op][k]+=float(r[ix[k]] or 0)
            except: pass
for op,(n,e) in sorted(byop.items(), key=lambda x:-x[1][0])[:10]:
    print("%-8s samples %6d (%.1f%%) executed %d"%(op,n,100*n/tot,e))
print({k:round(100*v/tot,1) for k,v in stall.items()})
for op,d in agg.items(): print(op, {k:int(v) for k,v in d.items()})
top=sorted(data,key=lambda r:-int(r[ix["# Samples"]] or 0))[:12]
for r in top:
    print(r[ix["Address"]][-5:], r[ix["# Samples"]], r[ix["Source"]][:60], "| mio",r[ix["stall_mio"]],"ssb",r[ix["stall_short_sb"]],"lsb",r[ix["stall_long_sb"]],"wait",r[ix["stall_wait"]],"math",r[ix["stall_math"]], "lg", r[ix["stall_lg"]])
