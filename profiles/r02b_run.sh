# usage: bash profiles/r02b_run.sh  -- GPU tests (whole suite, then the no-alloc test alone in a fresh process), benches, ncu
mkdir -p gpurun_out; TAG=r02b
(time timeout 1200 python -m pytest tests -m gpu -q -x) > gpurun_out/${TAG}_pytest.log 2>&1; tail -4 gpurun_out/${TAG}_pytest.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/${TAG}_pytest_noalloc_alone.log 2>&1; tail -2 gpurun_out/${TAG}_pytest_noalloc_alone.log
python bench.py --steps 100 --warmup 10 > gpurun_out/${TAG}_bench_chignolin.json 2> gpurun_out/${TAG}_bench_chignolin.err
B="python bench.py --workload water --steps 20 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/${TAG}_bench_water.json 2> gpurun_out/${TAG}_bench_water.err
TMDN_EDGE_PERSIST=0 $B > gpurun_out/${TAG}_bench_water_nopersist.json 2> gpurun_out/${TAG}_bench_water_nopersist.err
TMDN_EDGE_PERSIST=0 $B --table-intervals 4096 --no-table-warp > gpurun_out/${TAG}_bench_water_oldtable.json 2> gpurun_out/${TAG}_bench_water_oldtable.err
TMDN_HOT_ROWS=24 $B > gpurun_out/${TAG}_bench_water_hot24.json 2> gpurun_out/${TAG}_bench_water_hot24.err
for w in md17 qm9-et md17-et; do
  python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_$w.json 2> gpurun_out/${TAG}_bench_$w.err
done
ncu --set full --clock-control none --import-source on --kernel-name regex:"tn_msg|tn_embed_edge" --launch-skip 18 --launch-count 6 -o gpurun_out/${TAG}_full_water98k -f $B --steps 1 > gpurun_out/${TAG}_ncu_water.log 2>&1
python - <<PY
import json
for n in ["chignolin","water","water_nopersist","water_oldtable","water_hot24","md17","qm9-et","md17-et"]:
    try:
        d=json.load(open("gpurun_out/${TAG}_bench_%s.json"%n)); kf=d["kernel_families"]
        print(n, "ms", round(d["ms_per_step"],4), "val", round(d["value"]), "e2e", round(d["e2e"]["value"]), "L", d["launches_per_step"],
              {k: round(v["us_per_step"],1) for k,v in kf.items() if k.startswith(("tn_msg","tn_embed_edge","et_edge"))})
    except Exception as e: print(n, "FAILED", e)
PY
