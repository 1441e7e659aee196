cd $GRAFT_REPO_ROOT
nvidia-smi topo -m > gpurun_out/c20_topo.txt 2>&1
timeout 400 python tools/hostlink_bench.py --ranks 1,2,4,8 --out gpurun_out/hostlink_r02.json > gpurun_out/c20_hostlink.log 2>&1; echo "hostlink exit $?"; cat gpurun_out/c20_hostlink.log | tail -6
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/c20_bench8.json 2> gpurun_out/c20_bench8.err; echo "bench8 exit $?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/c20_bench4.json 2> gpurun_out/c20_bench4.err; echo "bench4 exit $?"
python - <<'PY'
import json
for f in ("gpurun_out/c20_bench8.json","gpurun_out/c20_bench4.json"):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d["n_gpus"], d["value"], "e2e", d["e2e"]["value"], "f32", d["e2e_f32"]["value"], "serial", d.get("e2e_serial",{}).get("value"), "sustained", d["sustained"]["value"])
    except Exception as e: print(f, "ERR", e)
PY
