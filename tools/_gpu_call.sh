cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_pipeline_gpu.py -m gpu -x -q > gpurun_out/c7_runner.log 2>&1; echo "exit $?"
tail -25 gpurun_out/c7_runner.log
