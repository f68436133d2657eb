# usage (2 GPUs): bash profiles/r02n_run.sh -- CTA count of the halo push kernel (1.03 M-atom box on 2 GPUs); cell-list kernel breakdown
mkdir -p gpurun_out; TAG=r02n
for c in 0 296 64 16; do
  TMDN_PUSH_CTAS=$c timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 \
    bench.py --gpus 2 --workload water --water-side 70 --steps 8 --warmup 3 > gpurun_out/${TAG}_water1m_2gpu_push$c.json 2> gpurun_out/${TAG}_water1m_2gpu_push$c.err
  echo "push $c rc=$?"
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/${TAG}_launches_nbr1m.csv python neighbor_sweep.py --sizes 1048576 --cutoffs 5.0 > gpurun_out/${TAG}_ncu_nbr.log 2>&1; echo "ncu rc=$?"
python - <<PY
import json
for c in (0,296,64,16):
    try:
        ss=json.load(open("gpurun_out/${TAG}_water1m_2gpu_push%d.json"%c))["strong_scaling"]
        print("push ctas",c,"ms",round(ss["ms_per_step"],2),"noex",round(ss["ms_per_step_without_exchange"],2),"halo_ms",round(ss["halo_ms_per_step"],2),"exposed",round(ss["halo_exposed_ms_per_step"],2), ss["halo_per_point_rank0"])
    except Exception as e: print(c,"FAILED",e)
PY
