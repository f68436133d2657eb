# usage: bash profiles/quick_ncu_only.sh TAG [ENV=VAL ...]  -- ncu launch list of two chignolin steps
mkdir -p gpurun_out; TAG=$1; shift
env "$@" ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name regex:"tn_|sgemm|tc_|csr_|scan_|cell_" -c 120 --csv --log-file gpurun_out/${TAG}_launches_chig.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu.log 2>&1
