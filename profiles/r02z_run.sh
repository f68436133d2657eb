# usage: bash profiles/r02z_run.sh -- A/B after FFMA-only forms in the adjoint message kernel
mkdir -p gpurun_out; TAG=r02z2
(time timeout 900 python -m pytest tests/test_gpu_tensornet.py -m gpu -q -x) > gpurun_out/${TAG}_pytest.log 2>&1; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log; grep -E "^FAILED|Error|assert" gpurun_out/${TAG}_pytest.log | head -8
for w in water md17; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_${w}.json 2> gpurun_out/${TAG}_bench_${w}.err
  python -c "
import json
d=json.load(open('gpurun_out/${TAG}_bench_${w}.json')); print('$w', round(d['ms_per_step'],3), {k:(round(v['us_per_step']/1000,3), v['launches_per_step']) for k,v in d['kernel_families'].items() if k.startswith('tn_msg') or k.startswith('tn_embed_edge')}, d['energy_checksum'])
"
done
