cd $GRAFT_REPO_ROOT
MDE_CONV_SMALL_ALL=1 timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "conv" 2>&1 | tail -2
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_forward_gpu.py tests/test_parity_bench_gpu.py tests/test_pipeline_gpu.py -x -q -m gpu 2>&1 | tail -3
