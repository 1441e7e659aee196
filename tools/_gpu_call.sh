cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_forward_gpu.py -x -q -m gpu 2>&1 | tail -2
for i in 1 2; do
timeout 300 python bench.py --no-extras --steps 30 --warmup 5 2>> gpurun_out/c83_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('patch embed through TMA', d['value'], d['ms_per_step'])"
MDE_PATCH_EMBED_DIRECT=1 timeout 300 python bench.py --no-extras --steps 30 --warmup 5 2>> gpurun_out/c83_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('direct stores', d['value'], d['ms_per_step'])"
done
