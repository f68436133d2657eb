# usage: bash profiles/r03c_run.sh -- ncu counters of tn_msg_bwd_p at 1.03 M atoms with and without the hot-row branch in the loop
mkdir -p gpurun_out; TAG=r03c
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_bytes.sum,l1tex__t_sector_hit_rate.pct,l1tex__m_xbar2l1tex_read_bytes.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,lts__throughput.avg.pct_of_peak_sustained_elapsed,dram__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed
for hot in 0 1; do
TMDN_HOT_ROWS=$hot timeout 900 ncu --metrics $M --clock-control none --kernel-name regex:"tn_msg_(bwd|fwd)_p" --launch-count 4 --csv --log-file gpurun_out/${TAG}_ncu_hot$hot.csv python bench.py --steps 1 --warmup 3 --strong-steps 1 --no-cpu-baseline > gpurun_out/${TAG}_ncu_hot$hot.log 2>&1; echo "ncu hot=$hot rc=$?"
done
