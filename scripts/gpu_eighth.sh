#!/bin/bash
mkdir -p gpurun_out
python scripts/eighth_call.py > gpurun_out/r02ap_eighth_call.log 2>&1
RTC_TIMING=1 python scripts/eighth_call.py 2>&1 | grep "rank 0 of 8" | tail -4 >> gpurun_out/r02ap_eighth_call.log
cat gpurun_out/r02ap_eighth_call.log
