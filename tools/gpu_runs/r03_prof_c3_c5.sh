# round 3, 1 GPU: (1) the reference's full-solve trip count for C5 (matrix from the device generator) into the parity
# fixture; (2) ncu --set full of one SpMV(+p.Ap) launch of C3 (27-point, TMA tile engine) and of C5 (two SELL panel
# launches), each after the same command has exited 0 without ncu; raw pages exported next to the reports
mkdir -p gpurun_out
timeout 900 python tools/make_bench_parity.py --device --full random_spd_10m > gpurun_out/r03_c5_full_fixture.log 2>&1; tail -2 gpurun_out/r03_c5_full_fixture.log | cut -c1-400
cp tests/golden/bench_parity.json gpurun_out/r03_bench_parity_with_c5_full.json
for W in stencil3d_27pt_256 random_spd_10m; do
  CMD="python bench.py --workload $W --steps 2 --warmup 1 --configs none --no-e2e --no-cpu"
  timeout 600 $CMD > gpurun_out/r03_forprof_$W.json 2> gpurun_out/r03_forprof_$W.err || { echo "bench $W failed"; tail -c 800 gpurun_out/r03_forprof_$W.err; continue; }
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"^(cg_spmv_kernel|spmv_kernel)" -s 40 -c 2 -f -o gpurun_out/r03_prof_$W $CMD > gpurun_out/r03_ncu_$W.log 2>&1
  tail -2 gpurun_out/r03_ncu_$W.log
  ncu -i gpurun_out/r03_prof_$W.ncu-rep --page raw --csv > gpurun_out/r03_raw_$W.csv 2>/dev/null
  wc -c gpurun_out/r03_raw_$W.csv
done
ls -la gpurun_out/*.ncu-rep
