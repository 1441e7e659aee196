cd $GRAFT_REPO_ROOT
for i in 1 2; do
timeout 300 python bench.py --no-extras --steps 30 --warmup 5 2>> gpurun_out/c55_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('fused mlp', d['value'], d['ms_per_step'])"
MDE_FUSED_MLP=0 timeout 300 python bench.py --no-extras --steps 30 --warmup 5 2>> gpurun_out/c55_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('two gemms', d['value'], d['ms_per_step'])"
done
timeout 300 python -m pytest tests/test_forward_gpu.py -x -q -m gpu 2>&1 | tail -2
