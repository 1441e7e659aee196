cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "layernorm or vit" 2>&1 | tail -1
for i in 1 2; do
timeout 300 python bench.py --no-extras --steps 30 --warmup 5 2>> gpurun_out/c75_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('LN prefetch', d['value'], d['ms_per_step'])"
MDE_LIB_PATH=tools/_libmde_base2.so timeout 300 python bench.py --no-extras --steps 30 --warmup 5 2>> gpurun_out/c75_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('before', d['value'], d['ms_per_step'])"
done
