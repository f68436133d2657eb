# usage: bash profiles/r03i_run.sh -- filter-table size sweep on the 98 304-atom water box
mkdir -p gpurun_out; TAG=r03i
for n in 1024 768 512 384; do
  timeout 600 python bench.py --workload water --steps 20 --warmup 3 --no-cpu-baseline --table-intervals $n > gpurun_out/${TAG}_bench_water_t$n.json 2> gpurun_out/${TAG}_bench_water_t$n.err
  python -c "
import json
d=json.load(open('gpurun_out/${TAG}_bench_water_t$n.json')); print('$n', round(d['ms_per_step'],3), {k:(round(v['us_per_step']/1000,3)) for k,v in d['kernel_families'].items() if k.startswith('tn_msg') or k.startswith('tn_embed_edge')}, d['energy_checksum'])
"
done
