mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2x_launches_bench_step.csv python bench.py --steps 2 --warmup 1 --blocks parity --no-cpu-baseline > gpurun_out/r2x_launches.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/r2x_launches.log | cut -c1-300; wc -l gpurun_out/r2x_launches_bench_step.csv
