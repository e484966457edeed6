#!/bin/bash
mkdir -p gpurun_out
for spp in 125 1000; do timeout 120 python tools/timeline.py $spp; done 2>&1 | tee gpurun_out/r31_timeline.log
