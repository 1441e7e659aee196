cd $GRAFT_REPO_ROOT
for i in 1 2 3; do
timeout 300 python bench.py --no-extras --steps 30 --warmup 5 2>> gpurun_out/c82_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('range_finalize v2', d['value'], d['ms_per_step'])"
MDE_LIB_PATH=tools/_libmde_base2.so timeout 300 python bench.py --no-extras --steps 30 --warmup 5 2>> gpurun_out/c82_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('before', d['value'], d['ms_per_step'])"
done
