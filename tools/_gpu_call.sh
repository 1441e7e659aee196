cd $GRAFT_REPO_ROOT
( python tools/io_bench.py; MDE_LIB_PATH=tools/_libmde_scan16.so python tools/io_bench.py; python tools/io_bench.py 8 1024; python tools/io_bench.py 4 518 ) > gpurun_out/scan2.txt 2>&1; cat gpurun_out/scan2.txt
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py tests/test_forward_gpu.py -x -q -m gpu 2>&1 | tail -2
