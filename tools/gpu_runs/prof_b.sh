# 1 GPU: the bench command, then the ncu launch list of the same command (per-launch durations; shares, not absolutes)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --configs none --no-e2e --no-cpu"
timeout 600 $CMD > gpurun_out/bench_forprof.json 2> gpurun_out/bench_forprof.err || { echo "bench failed"; tail -c 800 gpurun_out/bench_forprof.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r02.csv $CMD > gpurun_out/ncu_launches.log 2>&1
tail -3 gpurun_out/ncu_launches.log; wc -l gpurun_out/launches_r02.csv
