# 8 GPUs: multi-GPU parity tests (log kept), the bench line, the timing traces, the launch-per-kernel loop for comparison
mkdir -p gpurun_out; TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 900 python -m pytest tests/test_dist_gpu.py -q -v -k "${CGB_N8_TESTS:-8-default or 8-jitter or 8-no_allgather or 4-default or 2-default}" > gpurun_out/pytest_dist_2_4_8.log 2>&1; tail -12 gpurun_out/pytest_dist_2_4_8.log
timeout 600 $TR bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; tail -c 300 gpurun_out/bench_n8.err
CGB_DIST_TIMING=1 timeout 300 $TR bench.py --gpus 8 --steps 1 --warmup 1 --configs none > gpurun_out/bench_n8_timing.json 2> gpurun_out/bench_n8_timing.err
grep "timing\]" gpurun_out/bench_n8_timing.err | tail -40 | cut -c1-330
CGB_DIST_PERSISTENT=0 timeout 300 $TR bench.py --gpus 8 --steps 2 --warmup 1 --configs none --no-e2e > gpurun_out/bench_n8_graph.json 2> gpurun_out/bench_n8_graph.err
python - <<PY
import json
for f in ("bench_n8","bench_n8_timing","bench_n8_graph"):
    try:
        d=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1])
        print(f, round(d["value"],1), d["run"]["loop"][:30], d["parity"].get("ok"), (d.get("e2e") or {}).get("value"))
        for k,v in d["configs"].items(): print("  ",k,round(v["value"],1),v["run"]["loop"][:30],v["run"].get("exchange"),v["parity"].get("ok"),(v.get("e2e") or {}).get("value"))
    except Exception as e:
        print(f,"ERR",e); print(open("gpurun_out/%s.err"%f).read()[-1500:])
PY
