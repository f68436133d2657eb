# usage (8 GPUs): bash profiles/r02g_run.sh -- strong scaling of the 1.03 M-atom box at 8 and 4 GPUs, both transports, slab vs brick grid
mkdir -p gpurun_out; TAG=r02g
run() {  # run <n> <name> <extra args...>
  n=$1; name=$2; shift 2
  timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $n --steps 20 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$?"; grep -v "Warning\|warn\|^\*\*\*\|NCCL version" gpurun_out/${TAG}_$name.err | tail -3
}
run 8 bench_8gpu
run 8 bench_8gpu_nccl --transport nccl
run 8 bench_8gpu_slab --grid 8x1x1
run 4 bench_4gpu
python - <<PY
import json
for n in ["bench_8gpu","bench_8gpu_nccl","bench_8gpu_slab","bench_4gpu"]:
    try:
        d=json.load(open("gpurun_out/${TAG}_%s.json"%n)); ss=d.get("strong_scaling") or {}
        print(n, "weak: ms", round(d["ms_per_step"],4), "val", round(d["value"]), "| strong:", {k: ss.get(k) for k in ("ms_per_step","value","grid","transport","ghost_atoms_max","boundary_atoms_max","halo_bytes_per_step","halo_ms_per_step","halo_exposed_ms_per_step","ms_per_step_without_exchange","launches_per_step","transport_note")}, "corr", {k:v for k,v in (ss.get("correctness") or {}).items() if k in ("max_abs_dF","rel_dE")}, "frac", (ss.get("roofline") or {}).get("frac"), ss.get("halo_per_point_rank0"))
    except Exception as e: print(n, "FAILED", e)
PY
