#!/bin/bash
# round 2, pass b (1 GPU): drop-in acceptance (the reference's own tests), whole -m gpu suite
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_dropin_gpu.py -m gpu -x -q > gpurun_out/r02b_pytest_dropin.log 2>&1; echo "dropin rc=$?"; tail -40 gpurun_out/r02b_pytest_dropin.log | cut -c1-400
cat gpurun_out/dropin_package.json gpurun_out/dropin_extension.json 2>/dev/null
timeout 1500 python -m pytest tests -m gpu -q --deselect tests/test_dropin_gpu.py > gpurun_out/r02b_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02b_pytest_gpu.log | cut -c1-300
