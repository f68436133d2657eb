# usage: bash profiles/r02f_run.sh -- full GPU suite; brute sweep (hit masks, division-free minimum image); ET benches (shared sigmoid)
mkdir -p gpurun_out; TAG=r02f
(time timeout 1200 python -m pytest tests -m gpu -q -x) > gpurun_out/${TAG}_pytest.log 2>&1; tail -4 gpurun_out/${TAG}_pytest.log
timeout 600 python neighbor_sweep.py --cutoffs 5.0 --sizes 1000,4000,16000,32767,65536 > gpurun_out/${TAG}_neighbor_sweep_rc5.jsonl 2> gpurun_out/${TAG}_sweep.err; echo "sweep rc=$?"
for w in qm9-et md17-et; do
  python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_$w.json 2> gpurun_out/${TAG}_bench_$w.err
done
python - <<PY
import json
for l in open("gpurun_out/${TAG}_neighbor_sweep_rc5.jsonl"):
    d=json.loads(l)
    if d.get("op")=="OptimizedDistance" and d["strategy"]=="brute": print((d["n"],d["periodic"],d["strategy"]), "%.3f ms"%d["ms"])
for n in ["qm9-et","md17-et"]:
    try:
        d=json.load(open("gpurun_out/${TAG}_bench_%s.json"%n)); kf=d["kernel_families"]
        print(n,"ms",round(d["ms_per_step"],3),{k: round(x["us_per_step"],1) for k,x in kf.items() if k.startswith("et_edge")})
    except Exception as e: print(n,"FAILED",e)
PY
