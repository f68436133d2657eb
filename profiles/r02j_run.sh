# usage: bash profiles/r02j_run.sh -- GN + prior + deploy tests, then the whole suite
mkdir -p gpurun_out; TAG=r02j
(time timeout 600 python -m pytest tests/test_gpu_gn.py tests/test_gpu_deploy.py -m gpu -q) > gpurun_out/${TAG}_pytest_new.log 2>&1; tail -25 gpurun_out/${TAG}_pytest_new.log | cut -c1-220
(time timeout 1200 python -m pytest tests -m gpu -q -x) > gpurun_out/${TAG}_pytest.log 2>&1; tail -4 gpurun_out/${TAG}_pytest.log
