# round 3, 2 GPUs: the whole -m gpu suite (world-2 multi-GPU cases included) on the build with the ADVICE fixes
# (cache bound, per-device stager, option check across ranks, side job behind the primed TMA pipeline), then the C4
# line with the pack's side job on the producer lane (default) against inline-before-the-tiles (CGB_DIST_SIDE_PACK=0)
mkdir -p gpurun_out; TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r03_pytest_gpu_n2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r03_pytest_gpu_n2.log; tail -6 gpurun_out/r03_pytest_gpu_n2.log
B="bench.py --gpus 2 --steps 3 --warmup 3 --configs none --no-e2e"
run() { tag=$1; shift; env "$@" timeout 300 $TR $B > gpurun_out/r03_n2_$tag.json 2> gpurun_out/r03_n2_$tag.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r03_n2_$tag.json").read().strip().splitlines()[-1])
    print("$tag", round(d["value"],1), "it/s", d["run"]["loop"][:40], "parity", d["parity"].get("ok"))
except Exception as e:
    print("$tag ERR", e); print(open("gpurun_out/r03_n2_$tag.err").read()[-1200:])
PY
}
run side1 CGB_X=0
run inline1 CGB_DIST_SIDE_PACK=0
run side2 CGB_X=0
run inline2 CGB_DIST_SIDE_PACK=0
CGB_DIST_TIMING=2 timeout 300 $TR $B > gpurun_out/r03_n2_timing.json 2> gpurun_out/r03_n2_timing.err; grep "persistent loop" gpurun_out/r03_n2_timing.err | tail -2 | cut -c1-400
