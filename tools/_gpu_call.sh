cd $GRAFT_REPO_ROOT
for i in 1 2; do
python tools/gemm_bench.py 32 50 2>&1 | grep -v "per CTA\|lifetime" | tr '\n' ' '; echo
MDE_LIB_PATH=tools/_libmde_halfa.so python tools/gemm_bench.py 32 50 2>&1 | grep -v "per CTA\|lifetime" | tr '\n' ' '; echo "(half A bytes)"
done > gpurun_out/halfa.txt 2>&1; cat gpurun_out/halfa.txt
