# 8 GPUs, quick: the bench line and every rank's phase trace
mkdir -p gpurun_out; TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; tail -c 300 gpurun_out/bench_n8.err
CGB_DIST_TIMING=2 timeout 300 $TR bench.py --gpus 8 --steps 1 --warmup 1 --configs none --no-e2e > gpurun_out/bench_n8_timing.json 2> gpurun_out/bench_n8_timing.err
grep "persistent loop, 1798" gpurun_out/bench_n8_timing.err | tail -8 | cut -c1-360
python - <<PY
import json
for f in ("bench_n8","bench_n8_timing"):
    try:
        d=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1])
        print(f, round(d["value"],1), d["run"]["loop"][:30], d["parity"].get("ok"), (d.get("e2e") or {}).get("value"))
        for k,v in d["configs"].items(): print("  ",k,round(v["value"],1),v["run"]["loop"][:30],v["run"].get("exchange"),v["parity"].get("ok"),(v.get("e2e") or {}).get("value"))
    except Exception as e:
        print(f,"ERR",e); print(open("gpurun_out/%s.err"%f).read()[-1500:])
PY
