mkdir -p gpurun_out
L=gpurun_out/r2z_varrounds3.log
: > $L
run() { echo "== bench $*" >> $L; timeout 280 python bench.py --warmup 5 --steps 40 --blocks parity --no-cpu-baseline "$@" 2>>$L | python -c "
import sys,json
d=json.loads(sys.stdin.read().splitlines()[-1])
print('C2 device %.0f MPix/s %.2f ms | e2e %.0f %.2f ms | clocks %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['clocks']['sm_mhz']))
" >> $L 2>&1; }
for i in 1 2 3 4; do
run
run --debug-bits 128
done
cat $L
