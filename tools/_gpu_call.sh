cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/c5_noalloc_alone.log 2>&1; echo "alone exit $?"
tail -3 gpurun_out/c5_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c5_all.log 2>&1; echo "all exit $?"
tail -4 gpurun_out/c5_all.log
