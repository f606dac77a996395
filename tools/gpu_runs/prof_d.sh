# 1 GPU: L2 access-policy window A/B under ncu (DRAM bytes, L2 hit rate, duration per SpMV launch), C4 and C2, one ncu pass each arm set
mkdir -p gpurun_out
timeout 300 python tools/l2window_ncu.py 3d7 512 3 > gpurun_out/l2ncu_plain.log 2>&1 || { echo "probe failed"; tail -5 gpurun_out/l2ncu_plain.log; exit 1; }
M="dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_sectors_srcunit_tex_op_read.sum,gpu__time_duration.sum"
timeout 900 ncu --metrics $M --clock-control none -k regex:"cg_(spmv|init)_kernel" --csv --log-file gpurun_out/l2window_ncu_c4.csv python tools/l2window_ncu.py 3d7 512 3 > gpurun_out/l2ncu_c4.log 2>&1
grep ARM gpurun_out/l2ncu_c4.log; wc -l gpurun_out/l2window_ncu_c4.csv
