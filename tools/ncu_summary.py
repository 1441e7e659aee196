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

# --traffic <out.json> [M]: the capture holds one ViT layer (QKV, attention, out-proj, LN, fc1, fc2, LN — tools/profile.sh step 3);
# write the DRAM bytes per launch of its four GEMM classes where bench.py's roofline leg looks them up.
if '--traffic' in sys.argv:
    import json
    dst = sys.argv[sys.argv.index('--traffic') + 1]
    M = int(sys.argv[sys.argv.index('--traffic') + 2]) if len(sys.argv) > sys.argv.index('--traffic') + 2 else 32 * 197
    order = [(2304, 768, 'qkv'), (768, 768, 'proj'), (3072, 768, 'fc1'), (768, 3072, 'fc2')]
    gem = [r for r in rows[2:] if 'tc_gemm2_kernel' in r[kn]]
    ir, iw, it = hdr.index('dram__bytes_read.sum'), hdr.index('dram__bytes_write.sum'), hdr.index('gpu__time_duration.sum')
    scale = lambda i: {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[units[i]]
    ent = []
    for (N, K, name), r in zip(order, gem):
        ent.append({'layer': name, 'M': M, 'N': N, 'K': K, 'dram_bytes': round(float(r[ir]) * scale(ir) + float(r[iw]) * scale(iw)),
                    'dram_read_bytes': round(float(r[ir]) * scale(ir)), 'dram_write_bytes': round(float(r[iw]) * scale(iw)),
                    'ncu_duration_us': float(r[it]) * (1e-3 if units[it] == 'nsecond' else 1.0)})
    json.dump({'source': rep.split('/')[-1] + ' (ncu --set full --clock-control none, one launch each, cold L2)', 'gemm': ent}, open(dst, 'w'), indent=1)
    print('wrote', dst)
