# usage (2 GPUs): bash profiles/r03d_run.sh -- multi-process tests and the driver's scaling command at HEAD, N = 2
mkdir -p gpurun_out; TAG=r03d
(time timeout 600 python -m pytest tests/test_gpu_distributed.py -m gpu -q) > gpurun_out/${TAG}_pytest_dist.log 2>&1; grep -E "passed|failed" gpurun_out/${TAG}_pytest_dist.log
n=2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29521 \
    bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_${n}gpu.json 2> gpurun_out/${TAG}_bench_${n}gpu.err; echo "N=$n rc=$?"
python - <<PY
import json
n=2
d=json.load(open("gpurun_out/${TAG}_bench_%dgpu.json"%n)); ss=d["strong_scaling"]
print(n, "weak ms", round(d["ms_per_step"],4), "val", round(d["value"]), "| strong ms", round(ss["ms_per_step"],2), "val", round(ss["value"]), ss["grid"], ss["transport"], "exposed", round(ss["halo_exposed_ms_per_step"],2), "noex", round(ss["ms_per_step_without_exchange"],2), ss["correctness"]["max_abs_dF"], ss["correctness"]["rel_dE"], ss["clocks"])
PY
