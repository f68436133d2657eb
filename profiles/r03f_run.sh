# usage: bash profiles/r03f_run.sh -- ET / T / GN tests and ET workloads after the instruction diet of et_edge_bwd
mkdir -p gpurun_out; TAG=r03f
(time timeout 900 python -m pytest tests/test_gpu_et.py tests/test_gpu_gn.py tests/test_gpu_training.py tests/test_gpu_deploy.py -m gpu -q -x) > gpurun_out/${TAG}_pytest.log 2>&1; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log; grep -E "^FAILED|Error|assert" gpurun_out/${TAG}_pytest.log | head -8
for w in md17-et qm9-et; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_${w}.json 2> gpurun_out/${TAG}_bench_${w}.err
  python -c "
import json
d=json.load(open('gpurun_out/${TAG}_bench_${w}.json')); print('$w', round(d['ms_per_step'],3), round(d['value']), {k:(round(v['us_per_step']/1000,3), v['launches_per_step']) for k,v in d['kernel_families'].items() if k.startswith('et_edge')}, d['energy_checksum'])
"
done
