import csv, sys, subprocess
rep=sys.argv[1]
out=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=rows[0]; units=rows[1]
want=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__grid_size','launch__block_size','smsp__issue_active.avg.pct_of_peak_sustained_active','l1tex__throughput.avg.pct_of_peak_sustained_elapsed','lts__throughput.avg.pct_of_peak_sustained_elapsed','sm__throughput.avg.pct_of_peak_sustained_elapsed','launch__shared_mem_per_block_dynamic']
idx=[hdr.index(w) for w in want if w in hdr]
kn=hdr.index('Kernel Name')
print('kernel | '+' | '.join(hdr[i].replace('.pct_of_peak_sustained','%').replace('.avg','') for i in idx))
for r in rows[2:]:
    print(r[kn].split('(')[0][-40:]+' | '+' | '.join(r[i] for i in idx))
