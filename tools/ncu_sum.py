import csv,sys,subprocess,collections
rep=sys.argv[1]
raw=subprocess.run(["ncu","-i",rep,"--page","raw","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines()))
hdr=rows[0]; units=rows[1]
want=["Kernel Name","gpu__time_duration.sum","dram__bytes_read.sum","dram__bytes_write.sum","lts__t_bytes.sum","lts__t_sectors_op_read.sum","lts__throughput.avg.pct_of_peak_sustained_elapsed","l1tex__throughput.avg.pct_of_peak_sustained_elapsed","sm__warps_active.avg.pct_of_peak_sustained_active","launch__registers_per_thread","smsp__inst_executed.sum","sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active","smsp__issue_active.avg.pct_of_peak_sustained_active","l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum","sm__cycles_elapsed.max","lts__t_sector_hit_rate.pct"]
for w in want:
    for i,h in enumerate(hdr):
        if h==w: print(f"{h} [{units[i]}]:", [r[i][:40] for r in rows[2:]])
