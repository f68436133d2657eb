# usage: bash profiles/r02e_run.sh -- new tests (deploy, neighbours), brute sweep after the division-free minimum image, edge-kernel variants
mkdir -p gpurun_out; TAG=r02e
(time timeout 900 python -m pytest tests/test_gpu_deploy.py tests/test_gpu_neighbors.py tests/test_gpu_tensornet.py tests/test_zz_no_alloc.py -m gpu -q -x) > gpurun_out/${TAG}_pytest.log 2>&1; tail -6 gpurun_out/${TAG}_pytest.log
timeout 600 python neighbor_sweep.py --cutoffs 5.0 --sizes 1000,4000,16000,32767,65536,1048576 > gpurun_out/${TAG}_neighbor_sweep_rc5.jsonl 2> gpurun_out/${TAG}_sweep.err; echo "sweep rc=$?"
B="python bench.py --workload water --steps 20 --warmup 3 --no-cpu-baseline"
for v in 0 1 2 3; do
  TMDN_EDGE_VARIANT=$v $B > gpurun_out/${TAG}_bench_water_variant$v.json 2> gpurun_out/${TAG}_bench_water_variant$v.err
done
python - <<PY
import json
for l in open("gpurun_out/${TAG}_neighbor_sweep_rc5.jsonl"):
    d=json.loads(l)
    if d.get("op")=="OptimizedDistance": print((d["n"],d["periodic"],d["strategy"]), "%.3f ms"%d["ms"], "%.1f%%"%(100*d["frac_hbm"]))
for v in range(4):
    try:
        d=json.load(open("gpurun_out/${TAG}_bench_water_variant%d.json"%v)); kf=d["kernel_families"]
        print("variant",v,"ms",round(d["ms_per_step"],3),{k: round(x["us_per_step"],1) for k,x in kf.items() if k.startswith("tn_msg")})
    except Exception as e: print(v,"FAILED",e)
PY
