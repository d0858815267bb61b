mkdir -p gpurun_out
L=gpurun_out/r2g_n8.log
: > $L
nvidia-smi -L | wc -l >> $L
timeout 600 python -m pytest tests/test_nccl_gpu.py -x -q -m gpu 2>&1 | tail -5 >> $L
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2g_bench_n8.out 2> gpurun_out/r2g_bench_n8.err
echo "rc=$?" >> $L
tail -1 gpurun_out/r2g_bench_n8.out > gpurun_out/r2g_bench_line_n8.json
tail -5 gpurun_out/r2g_bench_n8.err >> $L
cat $L
