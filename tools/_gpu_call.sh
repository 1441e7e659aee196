cd $GRAFT_REPO_ROOT
for bn in 0 64 128 256; do
echo "BLOCK_N=$bn"
MDE_CONV_BLOCK_N=$bn python tools/conv_bench.py 256 256 14 15 32 50
MDE_CONV_BLOCK_N=$bn python tools/conv_bench.py 256 128 14 15 32 50
MDE_CONV_BLOCK_N=$bn python tools/conv_bench.py 256 128 28 30 32 50
MDE_CONV_BLOCK_N=$bn python tools/conv_bench.py 128 64 28 30 32 50
done > gpurun_out/convbn.txt 2>&1
cat gpurun_out/convbn.txt
