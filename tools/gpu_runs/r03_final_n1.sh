# round 3, 1 GPU: the committed state — whole -m gpu suite, smoke(), the default bench line (both arms)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r03_pytest_gpu_final.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r03_pytest_gpu_final.log; tail -4 gpurun_out/r03_pytest_gpu_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 1200 python bench.py > gpurun_out/bench_r03_n1.json 2> gpurun_out/bench_r03_n1.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_r03_n1.json").read().strip().splitlines()[-1])
print(round(d["value"],2), "it/s  e2e", round(d["e2e"]["value"],2), "frac", round(d["roofline"]["frac"],4), "parity", d["parity"]["ok"], d["parity"].get("full_solve"))
for k,v in d["configs"].items(): print("  ",k,round(v["value"],1), "frac", round(v["roofline"]["frac"],3) if v.get("roofline") else None, v["parity"].get("ok"), (v["parity"].get("full_solve") or {}).get("ok"), (v["parity"].get("full_solve") or {}).get("iterations"))
print("e2e_cpp", d.get("e2e_cpp"))
PY
