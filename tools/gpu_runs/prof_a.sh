# 1 GPU, no profiler: micro-benchmarks, the L2-window A/B with CUDA events, the whole GPU test suite, the bench line
mkdir -p gpurun_out
tools/_build/gather_bench 40 134217728 3 > gpurun_out/gather_bench.json 2>&1; tail -1 gpurun_out/gather_bench.json
timeout 300 python tools/h2d_probe.py 1 > gpurun_out/h2d_probe.txt 2>&1; tail -8 gpurun_out/h2d_probe.txt
timeout 300 examples/e2e_cpp 128 3 1 > gpurun_out/e2e_cpp.txt 2>&1; tail -3 gpurun_out/e2e_cpp.txt
: > gpurun_out/l2window.jsonl
for cfg in "3d7 512 128" "3d7 128 442" "3d27 256 128"; do
  timeout 200 python tools/l2window_probe.py $cfg default >> gpurun_out/l2window.jsonl 2>> gpurun_out/l2window.err
  CGB_L2_WINDOW=1 timeout 200 python tools/l2window_probe.py $cfg window_maxcarve >> gpurun_out/l2window.jsonl 2>> gpurun_out/l2window.err
  CGB_L2_WINDOW=1 CGB_L2_CARVE_MB=32 CGB_L2_HITRATIO=0.5 timeout 200 python tools/l2window_probe.py $cfg window_carve32_hit0.5 >> gpurun_out/l2window.jsonl 2>> gpurun_out/l2window.err
  CGB_L2_WINDOW=1 CGB_L2_WINDOW_MODE=k1 timeout 200 python tools/l2window_probe.py $cfg window_k1only >> gpurun_out/l2window.jsonl 2>> gpurun_out/l2window.err
done
cut -c1-260 gpurun_out/l2window.jsonl
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -c 400 gpurun_out/bench_n1.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n1.json").read().strip().splitlines()[-1])
    k=d["roofline"]["kernels"] if "kernels" in d["roofline"] else {}
    print("bench_n1", round(d["value"],1), d["roofline"].get("frac"), d["parity"].get("ok"), (d.get("e2e") or {}).get("value"), d.get("e2e_cpp"))
    for kk,v in d["configs"].items(): print("  ",kk,round(v["value"],1),v["parity"].get("ok"),(v.get("e2e") or {}).get("value"),(v["roofline"] or {}).get("l2"))
except Exception as e:
    print("ERR",e)
PY
