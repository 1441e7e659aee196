cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c13_all.log 2>&1; echo "all exit $?"
tail -4 gpurun_out/c13_all.log
timeout 300 python bench.py --no-extras --steps 30 --warmup 5 > gpurun_out/c13_bench.json 2> gpurun_out/c13_bench.err; cat gpurun_out/c13_bench.json | head -c 400
