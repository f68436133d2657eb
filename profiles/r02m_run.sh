# usage: bash profiles/r02m_run.sh -- occupancy variants of the persistent message kernels (98 k-atom water box), bit-identity check
mkdir -p gpurun_out; TAG=r02m
B="python bench.py --workload water --steps 20 --warmup 3 --no-cpu-baseline"
for v in 0 4 16 8 12 24; do
  TMDN_EDGE_VARIANT=$v $B > gpurun_out/${TAG}_bench_water_variant$v.json 2> gpurun_out/${TAG}_bench_water_variant$v.err
done
python - <<PY
import json
for v in (0,4,16,8,12,24):
    try:
        d=json.load(open("gpurun_out/${TAG}_bench_water_variant%d.json"%v)); kf=d["kernel_families"]
        print("variant",v,"ms",round(d["ms_per_step"],3),{k: round(x["us_per_step"],1) for k,x in kf.items() if k.startswith("tn_msg")}, "checksum", d["energy_checksum"])
    except Exception as e: print(v,"FAILED",e)
PY
