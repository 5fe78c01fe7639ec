import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[0]; units=rows[1]
want=["Kernel Name","gpu__time_duration.sum","sm__throughput.avg.pct_of_peak_sustained_elapsed","smsp__issue_active.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active","sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active","l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed","l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum","l1tex__data_pipe_lsu_wavefronts_mem_shared.sum","smsp__inst_executed.sum","sm__warps_active.avg.pct_of_peak_sustained_active","dram__bytes_read.sum","dram__bytes_write.sum","gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed","launch__registers_per_thread","launch__occupancy_limit_shared_mem","launch__grid_size","sm__cycles_elapsed.avg","smsp__cycles_active.avg"]
idx={w:hdr.index(w) for w in want if w in hdr}
stalls=[h for h in hdr if h.startswith("smsp__average_warp") and "issue_stalled" in h and h.endswith("_per_warp_active.pct")]
if not stalls: stalls=[h for h in hdr if "warp_issue_stalled" in h and "per_warp_active" in h]
print("missing",[w for w in want if w not in hdr])
for r in rows[2:]:
    print("----", r[idx["Kernel Name"]][:80])
    for w in want[1:]:
        if w in idx: print("   %-75s %s %s"%(w, r[idx[w]], units[idx[w]]))
    st=sorted([(float(r[hdr.index(s)] or 0), s) for s in stalls], reverse=True)[:5]
    for v,s in st: print("      stall %-60s %.1f"%(s.replace("smsp__average_warps_issue_stalled_","").replace("smsp__average_warp_latency_issue_stalled_",""), v))
