cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_terrain_gpu.py -m gpu -x -q -k "holes or guided" > gpurun_out/c6_filters.log 2>&1; echo "exit $?"
tail -25 gpurun_out/c6_filters.log
