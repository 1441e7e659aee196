cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_backward_gpu.py -m gpu -x -q -s -k "training_step or graph_captured or inference_after or learning_rate" > gpurun_out/c21_train.log 2>&1; echo "exit $?"
grep -E "worst running|passed|failed|Error|assert" gpurun_out/c21_train.log | tail -12
timeout 300 python bench.py --model train --no-extras > gpurun_out/c21_train_bench.json 2> gpurun_out/c21_train_bench.err; python -c "import json; d=json.loads(open('gpurun_out/c21_train_bench.json').read()); print('train', d['value'], d['ms_per_step'])"
