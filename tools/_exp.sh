mkdir -p gpurun_out
L=gpurun_out/r2i_thin_split2.log
: > $L
for dbg in 0 1 2 3 32 33 34 35; do
  echo "== A0 debug $dbg (1: no im2col build, 2: no stores, 32: unsplit epilogue)" >> $L
  timeout 120 python tools/conv_bench.py --layers A0 --debug $dbg 2>&1 | tail -1 >> $L
done
cat $L
