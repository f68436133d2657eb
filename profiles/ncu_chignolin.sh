# usage: bash profiles/ncu_chignolin.sh TAG  -- ncu launch list of the last two steps and one full capture of one step (166 atoms)
mkdir -p gpurun_out; TAG=$1
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name regex:"tn_|sgemm|tc_|csr_|scan_|cell_" -c 250 --csv --log-file gpurun_out/${TAG}_launches_chignolin.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name regex:"tn_|sgemm|tc_|csr_|scan_|cell_" --launch-skip 164 --launch-count 41 -o gpurun_out/${TAG}_full_chignolin -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu2.log 2>&1
ls -la gpurun_out/${TAG}_full_chignolin.ncu-rep
