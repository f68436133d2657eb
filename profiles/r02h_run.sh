# usage: bash profiles/r02h_run.sh -- full GPU suite at HEAD, smoke(), the driver's bench commands (default arm with strong_scaling, reference arm)
mkdir -p gpurun_out; TAG=r02h
(time timeout 1200 python -m pytest tests -m gpu -q -x) > gpurun_out/${TAG}_pytest.log 2>&1; tail -4 gpurun_out/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${TAG}_smoke.log
(time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 3) > gpurun_out/${TAG}_bench_default.json 2> gpurun_out/${TAG}_bench_default.err; echo "bench rc=$?"; tail -4 gpurun_out/${TAG}_bench_default.err
(time timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3) > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "ref rc=$?"; tail -4 gpurun_out/${TAG}_bench_reference.err
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_default.json")); ss=d["strong_scaling"]
print("default: ms", d["ms_per_step"], "value", d["value"], "e2e", d["e2e"]["value"], "launches", d["launches_per_step"], "roofline", d["roofline"]["frac"], d["roofline"]["traffic"], "cpu", d["cpu_baseline"])
print("strong:", ss["ms_per_step"], ss["value"], (ss.get("roofline") or {}).get("frac"))
r=json.load(open("gpurun_out/${TAG}_bench_reference.json")); print("reference:", r["steps"], r["warmup"], r["ms_per_step"], r["value"], r["cpu_baseline"]["cores"])
PY
