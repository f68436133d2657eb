# usage (2 GPUs): bash profiles/r02c_run.sh -- decomposition tests, decomposed water boxes with both transports, strong-scaling arm
mkdir -p gpurun_out; TAG=r02c
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
(time timeout 600 python -m pytest tests/test_gpu_distributed.py -m gpu -q -x) > gpurun_out/${TAG}_pytest_dist.log 2>&1; tail -15 gpurun_out/${TAG}_pytest_dist.log
for t in peer nccl; do
  timeout 300 $TR bench.py --gpus 2 --workload water --water-side 32 --steps 10 --warmup 3 --transport $t > gpurun_out/${TAG}_water98k_2gpu_$t.json 2> gpurun_out/${TAG}_water98k_2gpu_$t.err; echo "98k $t rc=$?"; tail -3 gpurun_out/${TAG}_water98k_2gpu_$t.err
done
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_2gpu.json 2> gpurun_out/${TAG}_bench_2gpu.err; echo "2gpu default rc=$?"; tail -3 gpurun_out/${TAG}_bench_2gpu.err
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline --transport nccl > gpurun_out/${TAG}_bench_2gpu_nccl.json 2> gpurun_out/${TAG}_bench_2gpu_nccl.err; echo "2gpu nccl rc=$?"; tail -3 gpurun_out/${TAG}_bench_2gpu_nccl.err
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_1gpu.json 2> gpurun_out/${TAG}_bench_1gpu.err; echo "1gpu rc=$?"; tail -3 gpurun_out/${TAG}_bench_1gpu.err
python - <<PY
import json
for n in ["water98k_2gpu_peer","water98k_2gpu_nccl","bench_2gpu","bench_2gpu_nccl","bench_1gpu"]:
    try:
        d=json.load(open("gpurun_out/${TAG}_%s.json"%n)); ss=d.get("strong_scaling") or {}
        print(n, "ms", round(d["ms_per_step"],4), "val", round(d["value"]), "| strong:", {k: ss.get(k) for k in ("atoms_total","ms_per_step","value","ghost_atoms_max","halo_bytes_per_step","halo_ms_per_step","ms_per_step_without_exchange","halo_exposed_ms_per_step","launches_per_step","transport","transport_note","grid","correctness")}, (ss.get("roofline") or {}).get("frac"))
    except Exception as e: print(n, "FAILED", e)
PY
