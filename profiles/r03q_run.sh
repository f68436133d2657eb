# usage: bash profiles/r03q_run.sh -- HEAD: full GPU suite, smoke, the driver's bench commands (default arm incl. strong_scaling at N = 1, reference arm)
mkdir -p gpurun_out; TAG=r03q
(time timeout 1500 python -m pytest tests -m gpu -q) > gpurun_out/${TAG}_pytest.log 2>&1; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log; grep -E "^FAILED|Error" gpurun_out/${TAG}_pytest.log | head -5
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${TAG}_smoke.log
timeout 900 python bench.py --gpus 1 --steps 100 --warmup 10 > gpurun_out/${TAG}_bench_default.json 2> gpurun_out/${TAG}_bench_default.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "ref rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_default.json")); ss=d["strong_scaling"]
print("default: ms", round(d["ms_per_step"],4), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "L", d["launches_per_step"], "frac", d["roofline"]["frac"], "traffic", d["roofline"]["traffic"], "alg", d["roofline"]["algorithmic_bytes_per_launch"], "cpu", d["cpu_baseline"]["value"])
print("strong N=1:", ss["ms_per_step"], ss["value"], (ss.get("roofline") or {}).get("frac"))
r=json.load(open("gpurun_out/${TAG}_bench_reference.json")); print("reference:", r["steps"], r["warmup"], r["ms_per_step"], r["value"])
PY
for w in water md17; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_${w}.json 2> gpurun_out/${TAG}_bench_${w}.err
  python -c "
import json
d=json.load(open('gpurun_out/${TAG}_bench_${w}.json')); print('$w', round(d['ms_per_step'],3), round(d['value']), {k:(round(v['us_per_step']/1000,3)) for k,v in d['kernel_families'].items() if k.startswith('tn_msg')}, d['energy_checksum'])
"
done
