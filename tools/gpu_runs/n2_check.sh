# 2 GPUs: multi-GPU parity tests (persistent loop, jitter stress, dictionary format) and the C4 / C2 lines with the phase trace
mkdir -p gpurun_out; TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 100 python tools/dist_single_probe.py 48 2>&1 | tail -2
timeout 700 python -m pytest tests/test_dist_gpu.py -x -q -k "2-default or 2-jitter or 2-dict or 2-no_allgather" > gpurun_out/pytest_dist2.log 2>&1; tail -3 gpurun_out/pytest_dist2.log
CGB_DIST_TIMING=1 timeout 300 $TR bench.py --gpus 2 --steps 3 --warmup 3 --configs poisson3d_7pt_128 --no-e2e > gpurun_out/b2_side.json 2> gpurun_out/b2_side.err
grep "persistent loop" gpurun_out/b2_side.err | sed -n '2p;$p' | cut -c1-330
CGB_DIST_OWN_HALO=0 timeout 300 $TR bench.py --gpus 2 --steps 3 --warmup 3 --configs poisson3d_7pt_128 --no-e2e > gpurun_out/b2_bar.json 2> gpurun_out/b2_bar.err
timeout 300 $TR bench.py --gpus 2 --steps 3 --warmup 3 --configs poisson3d_7pt_128 --no-e2e > gpurun_out/b2_own.json 2> gpurun_out/b2_own.err
python - <<PY
import json
for f in ("b2_side","b2_bar","b2_own"):
    try:
        d=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1])
        print(f, round(d["value"],1), d["run"]["loop"][:30], d["parity"].get("ok"))
        for k,v in d["configs"].items(): print("  ",k,round(v["value"],1),v["run"]["loop"][:40],v["parity"].get("ok"))
    except Exception as e:
        print(f,"ERR",e); print(open("gpurun_out/%s.err"%f).read()[-1500:])
PY
