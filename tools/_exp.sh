mkdir -p gpurun_out
L=gpurun_out/r2r_subpixel2.log
: > $L
timeout 900 python -m pytest tests/test_layers_gpu.py -x -q -m gpu -k "sub_pixel or pipelined" 2>&1 | tail -15 >> $L
timeout 900 python -m pytest tests/test_train_gpu.py tests/test_layers_gpu.py -x -q -m gpu 2>&1 | tail -3 >> $L
timeout 400 python bench.py --steps 20 --warmup 5 --blocks c1,c5 --no-cpu-baseline 2>>$L | python -c "
import sys,json
d=json.loads(sys.stdin.read().splitlines()[-1])
print('C2 %.0f MPix/s %.2f ms' % (d['value'], d['ms_per_step']))
print('c1', json.dumps(d['c1'])[:700])
print('c5', json.dumps(d['c5'])[:400])
" >> $L 2>&1
cat $L
