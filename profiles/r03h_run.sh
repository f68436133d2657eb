# usage: bash profiles/r03h_run.sh -- scripted models with priors
mkdir -p gpurun_out; TAG=r03h
(time timeout 900 python -m pytest tests/test_gpu_deploy.py -m gpu -q -x) > gpurun_out/${TAG}_pytest.log 2>&1; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log; grep -E "^FAILED|^E  " gpurun_out/${TAG}_pytest.log | head -12
