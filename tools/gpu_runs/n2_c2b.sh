# 2 GPUs: C2 as a config behind the C4 headline (what the default bench does), per-solve device times
mkdir -p gpurun_out; TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
B="bench.py --gpus 2 --steps 3 --warmup 3 --configs poisson3d_7pt_128 --no-e2e"
run() { tag=$1; shift; env CGB_BENCH_TRACE=1 "$@" timeout 300 $TR $B > gpurun_out/c2b_$tag.json 2> gpurun_out/c2b_$tag.err; grep 'bench trace' gpurun_out/c2b_$tag.err | cut -c1-200 | sed "s/^/$tag: /"; pgrep -a nvidia-smi | head -3; }
run a
run b
run c
B="bench.py --gpus 2 --steps 12 --warmup 3 --workload poisson3d_7pt_128 --configs none --no-e2e"
run headline_sampled
run headline_nosampler CGB_BENCH_NO_CLOCKS=1
run headline_sampled2
