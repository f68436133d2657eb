# usage: bash profiles/r02r_run.sh -- HEAD: full GPU suite, smoke, driver bench commands, all workloads, ncu of the ET edge kernels
mkdir -p gpurun_out; TAG=r02r
(time timeout 1200 python -m pytest tests -m gpu -q -x) > gpurun_out/${TAG}_pytest.log 2>&1; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${TAG}_smoke.log
timeout 900 python bench.py --gpus 1 --steps 100 --warmup 10 > gpurun_out/${TAG}_bench_default.json 2> gpurun_out/${TAG}_bench_default.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "ref rc=$?"
for w in md17 water qm9-et md17-et; do
  python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_$w.json 2> gpurun_out/${TAG}_bench_$w.err
done
ncu --set full --clock-control none --import-source on --kernel-name regex:"et_edge" --launch-skip 24 --launch-count 4 -o gpurun_out/${TAG}_full_md17et -f python bench.py --workload md17-et --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_et.log 2>&1; echo "ncu rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_default.json")); ss=d["strong_scaling"]
print("default: ms", round(d["ms_per_step"],4), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "L", d["launches_per_step"], "frac", d["roofline"]["frac"], "traffic", d["roofline"]["traffic"], "alg", d["roofline"]["algorithmic_bytes_per_launch"])
print("strong N=1:", ss["ms_per_step"], ss["value"], (ss.get("roofline") or {}).get("frac"))
r=json.load(open("gpurun_out/${TAG}_bench_reference.json")); print("reference:", r["steps"], r["warmup"], r["ms_per_step"], r["value"])
for n in ["md17","water","qm9-et","md17-et"]:
    d=json.load(open("gpurun_out/${TAG}_bench_%s.json"%n)); print(n, "ms", round(d["ms_per_step"],3), "value", round(d["value"]), "frac", round(d["roofline"]["frac"],4), d["roofline"]["kernel"])
PY
