set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/c1_smi.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c1_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/c1_tests.log
tail -5 gpurun_out/c1_tests.log
timeout 600 python bench.py > gpurun_out/c1_bench.json 2> gpurun_out/c1_bench.err; echo "bench exit $?"
timeout 300 python tools/hostlink_bench.py --ranks 1 --out gpurun_out/c1_hostlink.json > gpurun_out/c1_hostlink.log 2>&1
timeout 300 python bench.py --model enhanced --batch 4 --no-extras > gpurun_out/c1_enh_b4.json 2> gpurun_out/c1_enh_b4.err
timeout 300 python bench.py --model config4 --batch 2 --no-extras > gpurun_out/c1_cfg4_b2.json 2> gpurun_out/c1_cfg4_b2.err
timeout 300 python bench.py --model config4 --no-extras --steps 5 > gpurun_out/c1_cfg4_b16.json 2> gpurun_out/c1_cfg4_b16.err
timeout 400 python bench.py --model train > gpurun_out/c1_train.json 2> gpurun_out/c1_train.err
tail -3 gpurun_out/c1_*.err
cat gpurun_out/c1_bench.json | head -c 1500
