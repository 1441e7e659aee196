cd $GRAFT_REPO_ROOT
python tools/tail_bench.py 32 20 fused
cp monodepth-estimation_b200/libmde.so /tmp/plain.so
MDE_BUILD_PROF=1 python monodepth-estimation_b200/build.py --force > /dev/null 2>&1
python tools/tail_bench.py 32 1 fused 2>&1 | tail -9
cp /tmp/plain.so monodepth-estimation_b200/libmde.so
