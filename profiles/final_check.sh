# usage: bash profiles/final_check.sh TAG  -- what the driver runs at round end (+ every GPU test file on its own)
mkdir -p gpurun_out; TAG=$1
(time timeout 900 python -m pytest tests -x -q -m gpu) > gpurun_out/${TAG}_pytest.log 2>&1; grep "passed\|failed" gpurun_out/${TAG}_pytest.log
for f in tests/test_gpu_*.py; do echo "== $f"; timeout 600 python -m pytest $f -x -q -m gpu 2>&1 | tail -1; done > gpurun_out/${TAG}_pytest_per_file.log 2>&1; cat gpurun_out/${TAG}_pytest_per_file.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -3 gpurun_out/${TAG}_smoke.log
python bench.py > gpurun_out/${TAG}_bench_default.json 2> gpurun_out/${TAG}_bench_default.err; wc -l gpurun_out/${TAG}_bench_default.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2>/dev/null; wc -l gpurun_out/${TAG}_bench_reference.json
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_default.json")); print("default", d["ms_per_step"], d["value"], d["e2e"]["value"], d["launches_per_step"], d["clocks"], d["roofline"]["frac"], d["cpu_baseline"])
PY
