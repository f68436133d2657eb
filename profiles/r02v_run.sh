# usage: bash profiles/r02v_run.sh -- D2 prior, cluster CSR build for small systems, n-ary molecule search: tests, A/B of the 166-atom step
mkdir -p gpurun_out; TAG=r02v
(time timeout 1500 python -m pytest tests -m gpu -q) > gpurun_out/${TAG}_pytest.log 2>&1; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log; grep -E "^FAILED|Error" gpurun_out/${TAG}_pytest.log | head -5
ab() { # name, env...
  n=$1; shift
  env "$@" timeout 600 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-strong > gpurun_out/${TAG}_bench_chig_$n.json 2> gpurun_out/${TAG}_bench_chig_$n.err
  python -c "
import json
d=json.load(open('gpurun_out/${TAG}_bench_chig_$n.json')); print('$n', round(d['ms_per_step'],4), round(d['value']), round(d['e2e']['value']), d['launches_per_step'], {k:round(v['us_per_step'],1) for k,v in d['kernel_families'].items() if k in ('node_gemm_small','neighbors_csr','tn_mol_reduce')}, d['energy_checksum'])
"
}
ab new TMDN_SMALL_CSR=1
ab csr4 TMDN_SMALL_CSR=0
ab new2 TMDN_SMALL_CSR=1
ab csr4b TMDN_SMALL_CSR=0
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name regex:"tn_|sgemm|tc_|csr_|scan_|cell_" -c 200 --csv --log-file gpurun_out/${TAG}_launches_chig.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-strong > gpurun_out/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
