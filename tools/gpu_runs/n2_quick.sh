# 2 GPUs, quick: persistent-loop parity tests (default + jitter stress) and the C4 / C2 lines with every rank's phase trace
mkdir -p gpurun_out; TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 400 python -m pytest tests/test_dist_gpu.py -x -q -k "2-default or 2-jitter" > gpurun_out/pytest_dist2q.log 2>&1; tail -3 gpurun_out/pytest_dist2q.log
CGB_DIST_TIMING=2 CGB_BENCH_TRACE=1 timeout 300 $TR bench.py --gpus 2 --steps 2 --warmup 2 --configs poisson3d_7pt_128 --no-e2e > gpurun_out/b2q.json 2> gpurun_out/b2q.err
grep "persistent loop, 1798\|persistent loop, 442\|bench trace" gpurun_out/b2q.err | tail -6 | cut -c1-360
python -c "
import json;d=json.loads(open('gpurun_out/b2q.json').read().strip().splitlines()[-1]);print('N=2',round(d['value'],1),d['parity'].get('ok'))
for k,v in d['configs'].items(): print('  ',k,round(v['value'],1),v['parity'].get('ok'))"
