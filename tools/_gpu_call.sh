cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 600 python bench.py > gpurun_out/r02h_bench.json 2> gpurun_out/r02h_bench_err.txt; echo bench rc=$?
timeout 300 python bench.py --model train > gpurun_out/r02h_train_bench.json 2>> gpurun_out/r02h_bench_err.txt; echo train rc=$?
timeout 300 python bench.py --model config4 > gpurun_out/r02h_config4_b16.json 2>> gpurun_out/r02h_bench_err.txt; echo c4 rc=$?
timeout 300 python bench.py --model enhanced --no-extras > gpurun_out/r02h_enhanced.json 2>> gpurun_out/r02h_bench_err.txt; echo enh rc=$?
timeout 300 python bench.py --model enhanced --batch 4 --no-extras > gpurun_out/r02h_enhanced_b4.json 2>> gpurun_out/r02h_bench_err.txt; echo b4 rc=$?
timeout 300 python bench.py --batch 4 --no-extras > gpurun_out/r02h_standard_b4.json 2>> gpurun_out/r02h_bench_err.txt; echo sb4 rc=$?
timeout 300 python bench.py --batch 1 --size 518 --no-extras > gpurun_out/r02h_standard_b1_518.json 2>> gpurun_out/r02h_bench_err.txt; echo b1 rc=$?
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02h_reference.json 2>> gpurun_out/r02h_bench_err.txt; echo ref rc=$?
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
