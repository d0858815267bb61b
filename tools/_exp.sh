mkdir -p gpurun_out
for i in 1 2; do
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().splitlines()[-1])
print('run $i: C2 device %.0f MPix/s %.2f ms | e2e %.0f %.2f ms | clocks %s | c3 %.0f/%.0f | c5 %.2f | parity %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['clocks'], d['c3']['value'], d['c3']['e2e']['value'], d['c5']['ms_per_step'], d['parity']['strings_equal']))
"
done
