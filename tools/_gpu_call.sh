cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 600 python bench.py > gpurun_out/r02i_bench.json 2> gpurun_out/r02i_bench_err.txt; echo bench rc=$?
timeout 300 python bench.py --model enhanced --no-extras > gpurun_out/r02i_enhanced.json 2>> gpurun_out/r02i_bench_err.txt; echo enh rc=$?
timeout 300 python bench.py --model config4 --no-extras > gpurun_out/r02i_config4_b16.json 2>> gpurun_out/r02i_bench_err.txt; echo c4 rc=$?
timeout 300 python bench.py --model train --no-extras > gpurun_out/r02i_train_bench.json 2>> gpurun_out/r02i_bench_err.txt; echo train rc=$?
timeout 300 python bench.py --batch 4 --no-extras > gpurun_out/r02i_standard_b4.json 2>> gpurun_out/r02i_bench_err.txt; echo b4 rc=$?
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
