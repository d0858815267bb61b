mkdir -p gpurun_out
L=gpurun_out/r2q_train2.log
: > $L
timeout 900 python -m pytest tests/test_train_gpu.py -x -q -m gpu 2>&1 | tail -5 >> $L
timeout 300 python tools/train_bench.py >> $L 2>&1
timeout 300 python tools/train_profile.py --top 12 2>&1 | tail -15 >> $L
cat $L
