cd $GRAFT_REPO_ROOT
run() { name=$1; shift; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$? lines=$(wc -l < gpurun_out/$name.json)"; }
run r02h_standard_b4_8gpu --batch 4 --no-extras --steps 50 --warmup 5
run r02h_enhanced_b4_8gpu --model enhanced --batch 4 --no-extras --steps 50 --warmup 5
run r02h_config4_b2_8gpu --model config4 --batch 2 --no-extras --steps 10 --warmup 3
run r02h_train_8gpu --model train --steps 20 --warmup 5
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02h_*8gpu.json')):
    try:
        d=json.load(open(f)); print(f.split('/')[-1], d['metric'], d['value'], d['ms_per_step'], d['n_gpus'])
    except Exception as e: print(f, 'ERR', e)
PY
