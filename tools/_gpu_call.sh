cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r02l
timeout 200 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 | tee gpurun_out/r02l/pytest_gpu.txt
timeout 150 python bench.py > gpurun_out/r02l/bench.json 2> gpurun_out/r02l/bench_err.txt; echo "bench rc=$?"
for i in 1 2; do
timeout 60 python bench.py --no-extras --steps 30 --warmup 5 2>> gpurun_out/r02l/bench_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('patch embed through TMA', d['value'], d['ms_per_step'])"
MDE_PATCH_EMBED_DIRECT=1 timeout 60 python bench.py --no-extras --steps 30 --warmup 5 2>> gpurun_out/r02l/bench_err.txt | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('direct stores', d['value'], d['ms_per_step'])"
done
timeout 60 python bench.py --model enhanced --no-extras > gpurun_out/r02l/enhanced.json 2>> gpurun_out/r02l/bench_err.txt
timeout 60 python bench.py --model train --no-extras > gpurun_out/r02l/train.json 2>> gpurun_out/r02l/bench_err.txt
timeout 60 python bench.py --batch 4 --no-extras > gpurun_out/r02l/standard_b4.json 2>> gpurun_out/r02l/bench_err.txt
python -c "
import json
for f in ('bench','enhanced','train','standard_b4'):
    try:
        d=json.load(open('gpurun_out/r02l/%s.json'%f)); print(f, d['value'], d['ms_per_step'], d.get('gpu_launches'), (d.get('e2e') or {}).get('value'))
    except Exception as e: print(f, 'missing', e)
"
timeout 120 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 300 -c 400 --csv --log-file gpurun_out/r02l/launches.csv python bench.py --no-extras --steps 2 --warmup 3 > gpurun_out/r02l/ncu_stdout.txt 2>&1; echo "ncu rc=$?"
