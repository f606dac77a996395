# 1 GPU: the HBM mix probe, the whole GPU test suite, smoke(), the default bench line
mkdir -p gpurun_out
timeout 120 tools/_build/stream_bench 4 5 > gpurun_out/stream_bench.json 2>&1; cat gpurun_out/stream_bench.json
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -c 300 gpurun_out/bench_n1.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n1.json").read().strip().splitlines()[-1])
    r=d["roofline"]
    print("bench_n1", round(d["value"],1), r.get("frac"), r.get("stream_peaks"), r.get("frac_of_same_mix_stream"), d["parity"].get("ok"), (d.get("e2e") or {}).get("value"))
    for kk,v in d["configs"].items(): print("  ",kk,round(v["value"],1),v["parity"].get("ok"),(v.get("e2e") or {}).get("value"))
except Exception as e:
    print("ERR",e)
PY
