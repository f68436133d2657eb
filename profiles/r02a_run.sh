# usage: bash profiles/r02a_run.sh  -- round 2, first GPU pass: tests, table-grid / persistent-kernel A/B, ncu of the water box
mkdir -p gpurun_out; TAG=r02a
(time timeout 900 python -m pytest tests -m gpu -q -x) > gpurun_out/${TAG}_pytest.log 2>&1; tail -5 gpurun_out/${TAG}_pytest.log
python bench.py --steps 100 --warmup 10 > gpurun_out/${TAG}_bench_chignolin.json 2> gpurun_out/${TAG}_bench_chignolin.err
python bench.py --steps 100 --warmup 10 --no-cpu-baseline --table-intervals 4096 --no-table-warp > gpurun_out/${TAG}_bench_chignolin_oldtable.json 2> gpurun_out/${TAG}_bench_chignolin_oldtable.err
B="python bench.py --workload water --steps 20 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/${TAG}_bench_water.json 2> gpurun_out/${TAG}_bench_water.err
TMDN_EDGE_PERSIST=0 $B > gpurun_out/${TAG}_bench_water_nopersist.json 2> gpurun_out/${TAG}_bench_water_nopersist.err
TMDN_EDGE_PERSIST=0 $B --table-intervals 4096 --no-table-warp > gpurun_out/${TAG}_bench_water_oldtable.json 2> gpurun_out/${TAG}_bench_water_oldtable.err
TMDN_HOT_ROWS=0 $B > gpurun_out/${TAG}_bench_water_hot0.json 2> gpurun_out/${TAG}_bench_water_hot0.err
TMDN_HOT_ROWS=16 $B > gpurun_out/${TAG}_bench_water_hot16.json 2> gpurun_out/${TAG}_bench_water_hot16.err
TMDN_HOT_ROWS=32 $B > gpurun_out/${TAG}_bench_water_hot32.json 2> gpurun_out/${TAG}_bench_water_hot32.err
python bench.py --workload md17 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_md17.json 2> gpurun_out/${TAG}_bench_md17.err
ncu --set full --clock-control none --import-source on --kernel-name regex:"tn_msg|tn_embed_edge" --launch-skip 18 --launch-count 6 -o gpurun_out/${TAG}_full_water98k -f $B --steps 1 > gpurun_out/${TAG}_ncu_water.log 2>&1
ls -la gpurun_out/*.ncu-rep
python - <<PY
import json
for n in ["chignolin","chignolin_oldtable","water","water_nopersist","water_oldtable","water_hot0","water_hot16","water_hot32","md17"]:
    try:
        d=json.load(open("gpurun_out/${TAG}_bench_%s.json"%n)); kf=d["kernel_families"]
        print(n, "ms", round(d["ms_per_step"],4), "val", round(d["value"]), "e2e", round(d["e2e"]["value"]), "L", d["launches_per_step"],
              {k: round(v["us_per_step"],1) for k,v in kf.items() if k.startswith(("tn_msg","tn_embed_edge"))})
    except Exception as e: print(n, "FAILED", e)
PY
