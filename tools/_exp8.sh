mkdir -p gpurun_out
L=gpurun_out/r2v_n2.log
: > $L
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" >> $L 2>&1
timeout 600 python -m pytest tests/test_nccl_gpu.py -x -q -m gpu 2>&1 | tail -3 >> $L
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2v_bench_n2.out 2> gpurun_out/r2v_bench_n2.err
echo "rc=$?" >> $L
tail -1 gpurun_out/r2v_bench_n2.out > gpurun_out/r2v_bench_line_n2.json
python - >> $L <<'PY'
import json
d=json.loads(open('gpurun_out/r2v_bench_line_n2.json').read().strip().splitlines()[-1])
print('N=2 C2 device %.0f MPix/s %.2f ms | e2e %.0f | c3 %.0f / %.0f | c5 %.2f ms allreduce %.3f identical %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['c3']['value'], d['c3']['e2e']['value'], d['c5']['ms_per_step'], d['c5']['allreduce_ms'], d['c5']['replicas_identical']))
PY
cat $L
