# usage: bash profiles/r02l_run.sh -- training-mode tests, then the whole suite
mkdir -p gpurun_out; TAG=r02l
(time timeout 600 python -m pytest tests/test_gpu_training.py -m gpu -q) > gpurun_out/${TAG}_pytest_new.log 2>&1; grep -E "passed|failed|Error|assert|^E " gpurun_out/${TAG}_pytest_new.log | head -30 | cut -c1-250
(time timeout 1200 python -m pytest tests -m gpu -q -x) > gpurun_out/${TAG}_pytest.log 2>&1; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log
