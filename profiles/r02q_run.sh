mkdir -p gpurun_out; TAG=r02q
timeout 600 python neighbor_sweep.py --cutoffs 5.0 --sizes 65536,1048576,4194304 > gpurun_out/${TAG}_neighbor_sweep_rc5.jsonl 2> gpurun_out/${TAG}_sweep.err
timeout 600 python neighbor_sweep.py --cutoffs 10.0 --sizes 262144 >> gpurun_out/${TAG}_neighbor_sweep_rc5.jsonl 2>> gpurun_out/${TAG}_sweep.err
(timeout 600 python -m pytest tests/test_gpu_neighbors.py -m gpu -q -k "cell or golden or csr") > gpurun_out/${TAG}_pytest_nbr.log 2>&1; grep -E "passed|failed" gpurun_out/${TAG}_pytest_nbr.log
python - <<PY
import json
for l in open("gpurun_out/${TAG}_neighbor_sweep_rc5.jsonl"):
    d=json.loads(l)
    if "ms" in d and d["strategy"]=="cell": print(d["op"], (d["n"],d["cutoff"],d["periodic"]), "%.3f ms"%d["ms"], "%.1f%%"%(100*d["frac_hbm"]))
PY
