mkdir -p gpurun_out
L=gpurun_out/r2o_lds.log
: > $L
for dbg in 0 32; do
  echo "== A0 debug $dbg (32: unpipelined epilogue), reps 30, float32 input" >> $L
  timeout 120 python tools/conv_bench.py --layers A0 --debug $dbg --reps 30 --check 2>&1 | tail -1 >> $L
done
echo "== all layers, pair auto" >> $L
timeout 200 python tools/conv_bench.py --reps 10 2>&1 | tail -14 >> $L
timeout 900 python -m pytest tests/test_layers_gpu.py tests/test_train_gpu.py -x -q -m gpu 2>&1 | tail -3 >> $L
for dbg in 0 32; do
timeout 280 python bench.py --steps 20 --warmup 5 --blocks parity --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().splitlines()[-1])
print('C2 device %.0f MPix/s %.2f ms | e2e %.0f | parity %s | conv isolated %.2f ms' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['parity']['strings_equal'], d['roofline']['ms_per_step']))
" >> $L 2>&1
done
cat $L
