cd $GRAFT_REPO_ROOT
for d in 0 4 1 2 3 6; do MDE_HALO_DBG=$d python tools/conv_bench.py 16 16 224 240 32 50; done
for d in 0 4 1 2 3; do MDE_HALO_DBG=$d python tools/conv_bench.py 64 64 512 512 4 20; done
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_forward_gpu.py -x -q -m gpu 2>&1 | tail -3
for i in 1 2; do timeout 300 python bench.py --no-extras --steps 30 --warmup 5 2> gpurun_out/c30_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('bench', d['value'], d['ms_per_step'])"
MDE_LIB_PATH=tools/_libmde_base.so timeout 300 python bench.py --no-extras --steps 30 --warmup 5 2> gpurun_out/c30_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('bench base lib', d['value'], d['ms_per_step'])"; done
