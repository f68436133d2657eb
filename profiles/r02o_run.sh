# usage: bash profiles/r02o_run.sh -- fp64 / training tests, whole suite, neighbour sweep after the pad-fill change
mkdir -p gpurun_out; TAG=r02o
(time timeout 600 python -m pytest tests/test_gpu_training.py -m gpu -q) > gpurun_out/${TAG}_pytest_new.log 2>&1; grep -E "passed|failed|^E " gpurun_out/${TAG}_pytest_new.log | head -20 | cut -c1-250
(time timeout 1200 python -m pytest tests -m gpu -q -x) > gpurun_out/${TAG}_pytest.log 2>&1; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log
timeout 600 python neighbor_sweep.py --cutoffs 5.0 --sizes 1000,32767,65536,1048576,4194304 > gpurun_out/${TAG}_neighbor_sweep_rc5.jsonl 2> gpurun_out/${TAG}_sweep.err
python - <<PY
import json
for l in open("gpurun_out/${TAG}_neighbor_sweep_rc5.jsonl"):
    d=json.loads(l)
    if d.get("op")=="OptimizedDistance": print((d["n"],d["periodic"],d["strategy"]), "%.3f ms"%d["ms"], "%.1f%%"%(100*d["frac_hbm"]))
PY
