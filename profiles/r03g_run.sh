# usage: bash profiles/r03g_run.sh -- HEAD after the ET adjoint diet and the reciprocal-based sigmoid: full GPU suite, all workloads
mkdir -p gpurun_out; TAG=r03g
(time timeout 1500 python -m pytest tests -m gpu -q) > gpurun_out/${TAG}_pytest.log 2>&1; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log; grep -E "^FAILED" gpurun_out/${TAG}_pytest.log | head -5
for w in chignolin water md17 qm9-et md17-et; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline --no-strong > gpurun_out/${TAG}_bench_${w}.json 2> gpurun_out/${TAG}_bench_${w}.err
  python -c "
import json
d=json.load(open('gpurun_out/${TAG}_bench_${w}.json')); print('$w', round(d['ms_per_step'],4), round(d['value']), 'frac', round(d['roofline']['frac'],4), {k:(round(v['us_per_step']/1000,3), v['launches_per_step']) for k,v in d['kernel_families'].items() if k.startswith('tn_msg') or k.startswith('et_edge')}, d['energy_checksum'])
"
done
