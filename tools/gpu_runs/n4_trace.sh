# 4 GPUs: phase trace of every rank (edge ranks 0, 3; interior ranks 1, 2), C4
mkdir -p gpurun_out; TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511"
CGB_DIST_TIMING=2 CGB_BENCH_TRACE=1 timeout 300 $TR bench.py --gpus 4 --steps 2 --warmup 1 --configs none --no-e2e > gpurun_out/b4_trace.json 2> gpurun_out/b4_trace.err
grep "persistent loop, 1798\|bench trace" gpurun_out/b4_trace.err | tail -9 | cut -c1-360
python -c "import json;d=json.loads(open('gpurun_out/b4_trace.json').read().strip().splitlines()[-1]);print('N=4',round(d['value'],1),d['parity'].get('ok'))"
