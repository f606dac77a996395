# 2 GPUs: why C2 (128^3, latency regime) varies between runs — per-solve device times under a few switches
mkdir -p gpurun_out; TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
B="bench.py --gpus 2 --steps 6 --warmup 3 --workload poisson3d_7pt_128 --configs none --no-e2e"
run() { tag=$1; shift; env CGB_BENCH_TRACE=1 "$@" timeout 200 $TR $B > gpurun_out/c2_$tag.json 2> gpurun_out/c2_$tag.err; echo "$tag: $(grep 'bench trace' gpurun_out/c2_$tag.err | cut -c1-200) value=$(python -c "import json;print(round(json.loads(open('gpurun_out/c2_$tag.json').read().strip().splitlines()[-1])['value'],1))" 2>&1 | tail -1)"; }
run default1
run default2
run timing CGB_DIST_TIMING=1
run inlinepack CGB_DIST_SIDE_PACK=0
run barrierhalo CGB_DIST_OWN_HALO=0
run graph CGB_DIST_PERSISTENT=0
run default3
