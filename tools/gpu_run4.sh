#!/bin/bash
mkdir -p gpurun_out
python tools/diag_media.py > gpurun_out/r4_diag.log 2>&1; tail -40 gpurun_out/r4_diag.log
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r4_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r4_pytest.log
tail -12 gpurun_out/r4_pytest.log
