cd $GRAFT_REPO_ROOT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r02h_bench_8gpu.json 2> gpurun_out/r02h_bench_8gpu_err.txt; echo rc=$?
tail -c 600 gpurun_out/r02h_bench_8gpu.json
