# usage: bash profiles/r02d_run.sh -- full GPU test suite, neighbour sweep (this build vs the reference's CUDA kernels recompiled for sm_100)
mkdir -p gpurun_out; TAG=r02d
(time timeout 1200 python -m pytest tests -m gpu -q -x) > gpurun_out/${TAG}_pytest.log 2>&1; tail -4 gpurun_out/${TAG}_pytest.log
timeout 900 python neighbor_sweep.py --cutoffs 5.0 > gpurun_out/${TAG}_neighbor_sweep_rc5.jsonl 2> gpurun_out/${TAG}_sweep.err; echo "sweep rc=$?"
timeout 900 python neighbor_sweep.py --cutoffs 5.0 --impl reference > gpurun_out/${TAG}_neighbor_sweep_rc5_reference.jsonl 2> gpurun_out/${TAG}_sweep_ref.err; echo "ref sweep rc=$?"; tail -2 gpurun_out/${TAG}_sweep_ref.err
python - <<PY
import json
ours={}; ref={}
for l in open("gpurun_out/${TAG}_neighbor_sweep_rc5.jsonl"):
    d=json.loads(l)
    if d.get("op")=="OptimizedDistance": ours[(d["n"],d["periodic"],d["strategy"])]=d
for l in open("gpurun_out/${TAG}_neighbor_sweep_rc5_reference.jsonl"):
    d=json.loads(l)
    if "ms" in d: ref[(d["n"],d["periodic"],d["strategy"])]=d
for k in sorted(ours):
    o=ours[k]; r=ref.get(k); rs=ref.get((k[0],k[1],"shared"))
    print(k, "ours %.3f ms (%.1f%% hbm)"%(o["ms"],100*o["frac_hbm"]), "| ref %s"%("%.3f ms"%r["ms"] if r else "-"), "| ref shared %s"%("%.3f ms"%rs["ms"] if rs and k[2]=="brute" else ""), "pairs", o["pairs"], r["pairs"] if r else "")
PY
