# usage: bash profiles/r02i_run.sh -- deploy/prior tests at HEAD; ncu launch list + full capture of the 166-atom headline step
mkdir -p gpurun_out; TAG=r02i
(time timeout 900 python -m pytest tests/test_gpu_deploy.py tests/test_gpu_tensornet.py tests/test_gpu_et.py tests/test_zz_no_alloc.py -m gpu -q -x) > gpurun_out/${TAG}_pytest.log 2>&1; tail -4 gpurun_out/${TAG}_pytest.log
python bench.py --steps 100 --warmup 10 --no-strong > gpurun_out/${TAG}_bench_chignolin.json 2> gpurun_out/${TAG}_bench_chignolin.err; echo "bench rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name regex:"tn_|sgemm|tc_|csr_|scan_|cell_" -c 200 --csv --log-file gpurun_out/${TAG}_launches_chignolin.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-strong > gpurun_out/${TAG}_ncu1.log 2>&1; echo "ncu1 rc=$?"
ncu --set full --clock-control none --import-source on --kernel-name regex:"tn_|sgemm|tc_|csr_|scan_|cell_" --launch-skip 123 --launch-count 41 -o gpurun_out/${TAG}_full_chignolin -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-strong > gpurun_out/${TAG}_ncu2.log 2>&1; echo "ncu2 rc=$?"
ls -la gpurun_out/${TAG}_full_chignolin.ncu-rep
