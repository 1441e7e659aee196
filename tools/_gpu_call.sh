cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out/r02k
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 | tee gpurun_out/r02k/pytest_gpu.txt
timeout 600 python bench.py > gpurun_out/r02k/bench.json 2> gpurun_out/r02k/bench_err.txt; echo "bench rc=$?"
timeout 300 python bench.py --model enhanced --no-extras > gpurun_out/r02k/enhanced.json 2>> gpurun_out/r02k/bench_err.txt
timeout 300 python bench.py --model config4 --no-extras > gpurun_out/r02k/config4_b16.json 2>> gpurun_out/r02k/bench_err.txt
timeout 300 python bench.py --model train --no-extras > gpurun_out/r02k/train.json 2>> gpurun_out/r02k/bench_err.txt
timeout 300 python bench.py --batch 4 --no-extras > gpurun_out/r02k/standard_b4.json 2>> gpurun_out/r02k/bench_err.txt
timeout 300 python bench.py --batch 1 --size 518 --no-extras > gpurun_out/r02k/standard_b1_518.json 2>> gpurun_out/r02k/bench_err.txt
python -c "
import json
for f in ('bench','enhanced','config4_b16','train','standard_b4','standard_b1_518'):
    d=json.load(open('gpurun_out/r02k/%s.json'%f)); print(f, d['value'], d['ms_per_step'], d.get('gpu_launches'), (d.get('e2e') or {}).get('value'))
"
python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 300 -c 400 --csv --log-file gpurun_out/r02k/launches.csv python bench.py --no-extras --steps 2 --warmup 3 > gpurun_out/r02k/ncu_stdout.txt 2>&1; echo "ncu rc=$?"
