mkdir -p gpurun_out
L=gpurun_out/r3a_solo.log
: > $L
timeout 1200 python -m pytest tests/test_range_coder_gpu.py tests/test_oracle_golden.py tests/test_pipeline_gpu.py -x -q -m gpu 2>&1 | tail -4 >> $L
for dbg in 0 256; do
echo "== coder_parts --pack 8 --debug $dbg (256: chains replicated in 32 lanes)" >> $L
timeout 200 python tools/coder_parts.py --pack 8 --debug $dbg --reps 10 2>&1 | tail -4 >> $L
done
run() { echo "== bench $*" >> $L; timeout 280 python bench.py --warmup 5 --steps 40 --blocks parity --no-cpu-baseline "$@" 2>>$L | python -c "
import sys,json
d=json.loads(sys.stdin.read().splitlines()[-1])
print('C2 device %.0f MPix/s %.2f ms | e2e %.0f %.2f ms | clocks %s MHz %s W | parity %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['clocks']['sm_mhz'], d['clocks'].get('power_w'), d['parity']['strings_equal']))
" >> $L 2>&1; }
for i in 1 2 3 4; do
run
run --debug-bits 256
done
cat $L
