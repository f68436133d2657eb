# usage: bash profiles/r02k_run.sh -- GN / T tests, then the whole suite
mkdir -p gpurun_out; TAG=r02k
(time timeout 600 python -m pytest tests/test_gpu_gn.py -m gpu -q) > gpurun_out/${TAG}_pytest_new.log 2>&1; grep -E "passed|failed|Error|assert" gpurun_out/${TAG}_pytest_new.log | head -20 | cut -c1-220
