cd $GRAFT_REPO_ROOT
for sp in 1 2 3; do
  MDE_VIT_SPLIT=$sp timeout 300 python bench.py --no-extras --steps 30 --warmup 5 2> gpurun_out/c14_err_$sp.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('split', $sp, d['value'], d['ms_per_step'])"
done
MDE_VIT_SPLIT=2 timeout 600 python -m pytest tests/test_parity_bench_gpu.py tests/test_forward_gpu.py -m gpu -x -q 2>&1 | tail -3
