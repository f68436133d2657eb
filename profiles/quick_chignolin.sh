mkdir -p gpurun_out; TAG=$1
(time timeout 500 python -m pytest tests/test_gpu_tensornet.py tests/test_gpu_distributed.py -m gpu -q) > gpurun_out/${TAG}_pytest.log 2>&1; tail -3 gpurun_out/${TAG}_pytest.log
python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_bench_chig.json 2> gpurun_out/${TAG}_bench_chig.err
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name regex:"tn_|sgemm|tc_|csr_|scan_|cell_" -c 120 --csv --log-file gpurun_out/${TAG}_launches_chig.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu.log 2>&1
python -c "
import json
d=json.load(open('gpurun_out/${TAG}_bench_chig.json')); print('chig', d['ms_per_step'], d['value'], d['e2e']['value'], d['launches_per_step'])
"
