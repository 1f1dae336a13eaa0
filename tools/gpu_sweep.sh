#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/final_test.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final_test.log; tail -3 gpurun_out/final_test.log
{
for S in 40 80 110 148 296 1184; do python tools/perf_probe.py $S 240 2 1 0 2>&1 | tail -1; done
FVAD_TIMELINE=gpurun_out/timeline_592.csv python tools/perf_probe.py 592 60 2 1 1 2>&1 | head -2
} 2>&1 | tee gpurun_out/sweep.log
python tools/timeline_summary.py gpurun_out/timeline_592.csv > gpurun_out/timeline_592_summary.txt 2>&1
