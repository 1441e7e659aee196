cd $GRAFT_REPO_ROOT
timeout 80 python -m pytest tests/test_kernels_gpu.py tests/test_forward_gpu.py -x -q -m gpu 2>&1 | tail -1
timeout 40 python bench.py --no-extras --steps 30 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('final build', d['value'], d['ms_per_step'], d['gpu_launches'])"
