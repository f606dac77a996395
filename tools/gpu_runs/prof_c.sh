# 1 GPU: the bench command, then ONE ncu --set full capture of the three kernels of a trip (K3, K1, K2) of the same command
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --configs none --no-e2e --no-cpu"
timeout 600 $CMD > gpurun_out/bench_forprof.json 2> gpurun_out/bench_forprof.err || { echo "bench failed"; tail -c 800 gpurun_out/bench_forprof.err; exit 1; }
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"cg_(spmv|pupdate_tma|update_tma|pupdate|update)_kernel" -s 30 -c 3 -f -o gpurun_out/prof_r02_trip $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/*.ncu-rep
