# round 3, N >= 3 GPUs: C4 with the two edge ranks given extra rows (they have one halo plane instead of two and wait
# in the p.Ap all-reduce otherwise); CGB_BENCH_EDGE_PLANES = extra rows in planes.  usage: bash r03_edge_partition.sh N "0 1.25 2"
N=${1:-4}; SET=${2:-"0 1.25 2 0"}
mkdir -p gpurun_out; TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
B="bench.py --gpus $N --steps 3 --warmup 3 --configs none --no-e2e"
i=0
for E in $SET; do
  i=$((i+1)); tag="n${N}_edge${E}_$i"
  CGB_BENCH_EDGE_PLANES=$E timeout 400 $TR $B > gpurun_out/r03_$tag.json 2> gpurun_out/r03_$tag.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r03_$tag.json").read().strip().splitlines()[-1])
    print("$tag", round(d["value"],1), "it/s  frac/GPU", round(d["roofline"]["frac"],4), "parity", d["parity"].get("ok"), d["parity"].get("full_solve"), d["clocks"])
except Exception as e:
    print("$tag ERR", e); print(open("gpurun_out/r03_$tag.err").read()[-1200:])
PY
done
for E in 0 1.25; do
CGB_BENCH_EDGE_PLANES=$E CGB_DIST_TIMING=2 timeout 400 $TR $B > gpurun_out/r03_n${N}_timing_e$E.json 2> gpurun_out/r03_n${N}_timing_e$E.err
echo "edge=$E"; grep "persistent loop" gpurun_out/r03_n${N}_timing_e$E.err | tail -$N | cut -c1-330
done
