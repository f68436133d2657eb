# usage (8 GPUs): bash profiles/r03e_run.sh -- the driver's scaling command at HEAD, N = 8 and 4 (+ the reference arm under torchrun at N = 8)
mkdir -p gpurun_out; TAG=r03e
for n in 8 4; do
  timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29521 \
    bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_${n}gpu.json 2> gpurun_out/${TAG}_bench_${n}gpu.err; echo "N=$n rc=$?"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 \
    bench.py --impl reference --gpus 8 --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_reference_8gpu.json 2> /dev/null; echo "ref N=8 rc=$?"
python - <<PY
import json
for n in (8,4):
    d=json.load(open("gpurun_out/${TAG}_bench_%dgpu.json"%n)); ss=d["strong_scaling"]
    print(n, "weak ms", round(d["ms_per_step"],4), "val", round(d["value"]), "| strong ms", round(ss["ms_per_step"],2), "val", round(ss["value"]), ss["grid"], ss["transport"], "exposed", round(ss["halo_exposed_ms_per_step"],2), "noex", round(ss["ms_per_step_without_exchange"],2), ss["correctness"]["max_abs_dF"], ss["correctness"]["rel_dE"], ss["clocks"])
r=json.load(open("gpurun_out/${TAG}_bench_reference_8gpu.json")); print("   reference arm:", r["n_gpus"], r["steps"], round(r["value"],1))
PY
