# usage: bash profiles/r02x_run.sh -- small GEMM as 16x32 tiles / 2x2 per thread; cluster CSR; D2: full suite, 166-atom step, other workloads
mkdir -p gpurun_out; TAG=r02x
(time timeout 1500 python -m pytest tests -m gpu -q) > gpurun_out/${TAG}_pytest.log 2>&1; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log; grep -E "^FAILED|Error" gpurun_out/${TAG}_pytest.log | head -5
timeout 900 python bench.py --steps 100 --warmup 10 --no-strong > gpurun_out/${TAG}_bench_chignolin.json 2> gpurun_out/${TAG}_bench_chignolin.err
for w in md17 water qm9-et md17-et; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_$w.json 2> gpurun_out/${TAG}_bench_$w.err
done
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_chignolin.json"))
print("chignolin: ms", round(d["ms_per_step"],4), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "L", d["launches_per_step"], "frac", d["roofline"]["frac"], {k:round(v['us_per_step'],1) for k,v in d['kernel_families'].items() if k in ('node_gemm_small','neighbors_csr','tn_mol_reduce')}, d['energy_checksum'])
for n in ["md17","water","qm9-et","md17-et"]:
    d=json.load(open("gpurun_out/${TAG}_bench_%s.json"%n)); print(n, "ms", round(d["ms_per_step"],3), "value", round(d["value"]), "frac", round(d["roofline"]["frac"],4), d["roofline"]["kernel"])
PY
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name regex:"tn_|sgemm|tc_|csr_|scan_|cell_" -c 200 --csv --log-file gpurun_out/${TAG}_launches_chig.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-strong > gpurun_out/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
