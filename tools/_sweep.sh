for d in 0 2 6 18 16 22; do MDE_HALO_DBG=$d python tools/conv_bench.py 16 16 224 224 32 20; done
for d in 0 2 6 18; do MDE_HALO_DBG=$d python tools/conv_bench.py 16 1 224 224 32 20 1; done
