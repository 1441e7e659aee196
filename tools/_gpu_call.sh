cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_forward_gpu.py tests/test_parity_bench_gpu.py -x -q -m gpu 2>&1 | tail -3
timeout 300 python bench.py --model config4 --no-extras --steps 10 --warmup 3 2> gpurun_out/c33_err.txt | tee gpurun_out/c33_config4.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('config4', d['value'], d['ms_per_step'])"
MDE_LIB_PATH=tools/_libmde_base.so timeout 300 python bench.py --model config4 --no-extras --steps 10 --warmup 3 2>> gpurun_out/c33_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('config4 base lib', d['value'], d['ms_per_step'])"
timeout 300 python bench.py --model enhanced --no-extras --steps 20 --warmup 3 2>> gpurun_out/c33_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('enhanced', d['value'], d['ms_per_step'])"
MDE_LIB_PATH=tools/_libmde_base.so timeout 300 python bench.py --model enhanced --no-extras --steps 20 --warmup 3 2>> gpurun_out/c33_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('enhanced base lib', d['value'], d['ms_per_step'])"
