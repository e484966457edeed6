#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r11_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/r11_alone.log
tail -25 gpurun_out/r11_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r11_pytest.log 2>&1; echo "all rc=$?" >> gpurun_out/r11_pytest.log
tail -6 gpurun_out/r11_pytest.log
