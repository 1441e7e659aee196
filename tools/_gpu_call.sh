cd $GRAFT_REPO_ROOT
for b in 6 8 12 16; do
timeout 300 python bench.py --batch $b --no-extras --steps 50 --warmup 5 2>> gpurun_out/c37_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('batch $b pair', d['value'], d['ms_per_step'])"
MDE_GEMM_1CTA=1 timeout 300 python bench.py --batch $b --no-extras --steps 50 --warmup 5 2>> gpurun_out/c37_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('batch $b 1cta', d['value'], d['ms_per_step'])"
done
