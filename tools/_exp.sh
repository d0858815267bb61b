mkdir -p gpurun_out
L=gpurun_out/r2m_pmf.log
: > $L
timeout 900 python -m pytest tests/test_range_coder_gpu.py -x -q -m gpu -k "pmf" 2>&1 | tail -8 >> $L
timeout 600 python -m pytest tests/test_baseline_shapes_gpu.py -x -q -m gpu -k "sweep or 65536 or pmf" 2>&1 | tail -4 >> $L
timeout 400 python bench.py --steps 20 --warmup 5 --blocks c4 --no-cpu-baseline 2>>$L | python -c "
import sys,json
d=json.loads(sys.stdin.read().splitlines()[-1])
print('C2 %.0f MPix/s %.2f ms' % (d['value'], d['ms_per_step']))
for r in d['c4']['pmf_to_quantized_cdf']: print(r)
" >> $L 2>&1
cat $L
