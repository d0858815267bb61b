mkdir -p gpurun_out
L=gpurun_out/r2y_dual2.log
: > $L
for pk in 2 4 8; do
for dbg in 0 64; do
echo "== coder_parts --pack $pk --debug $dbg (64: one stream per chain warp)" >> $L
timeout 200 python tools/coder_parts.py --pack $pk --debug $dbg --reps 10 2>&1 | grep "y encode" >> $L
done; done
run() { echo "== bench $*" >> $L; timeout 280 python bench.py --warmup 5 --steps 20 --blocks parity --no-cpu-baseline "$@" 2>>$L | python -c "
import sys,json
d=json.loads(sys.stdin.read().splitlines()[-1])
print('C2 device %.0f MPix/s %.2f ms | e2e %.0f %.2f ms | parity %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['parity']['strings_equal']))
" >> $L 2>&1; }
run
run
run --debug-bits 64
run --debug-bits 64
cat $L
