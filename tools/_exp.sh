mkdir -p gpurun_out
L=gpurun_out/r2t_lanes_packed.log
: > $L
for cfg in "1 0" "2 64" "4 128" "8 256" "8 512" "16 512"; do
  set -- $cfg
  echo "== coder_parts --lanes 2 --repeat $1 --pack $2   (streams = 64 x $1)" >> $L
  timeout 300 python tools/coder_parts.py --lanes 2 --repeat $1 --pack $2 --reps 2 2>&1 | tail -4 >> $L
done
cat $L
