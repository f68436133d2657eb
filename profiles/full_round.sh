# usage: bash profiles/full_round.sh TAG   -- GPU tests, the five bench workloads, ncu launch list + full capture (chignolin)
mkdir -p gpurun_out; TAG=$1
(time timeout 700 python -m pytest tests -m gpu -q) > gpurun_out/${TAG}_pytest.log 2>&1; tail -3 gpurun_out/${TAG}_pytest.log | head -1
python bench.py --steps 100 --warmup 10 > gpurun_out/${TAG}_bench_chignolin.json 2> gpurun_out/${TAG}_bench_chignolin.err
for w in md17 water qm9-et md17-et; do
  python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_$w.json 2> gpurun_out/${TAG}_bench_$w.err
done
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference_arm.json 2> gpurun_out/${TAG}_bench_reference_arm.err
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name regex:"tn_|sgemm|tc_|csr_|scan_|cell_" -c 200 --csv --log-file gpurun_out/${TAG}_launches_chignolin.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name regex:"tn_|sgemm|tc_|csr_|scan_|cell_" --launch-skip 135 --launch-count 45 -o gpurun_out/${TAG}_full_chignolin -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu2.log 2>&1
ls -la gpurun_out/*.ncu-rep
python - <<PY
import json
for n in ["chignolin","md17","water","qm9-et","md17-et"]:
    try:
        d=json.load(open("gpurun_out/${TAG}_bench_%s.json"%n)); print(n, d["ms_per_step"], d["value"], d["e2e"]["value"], d["launches_per_step"])
    except Exception as e: print(n, "FAILED", e)
PY
