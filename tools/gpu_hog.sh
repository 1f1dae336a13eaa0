#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_golden.py -x -q -m gpu 2>&1 | tail -2
{
for S in 20 40 80 110; do python tools/perf_probe.py $S 600 2 1 0 2>&1 | tail -1; done
FVAD_TIMELINE=gpurun_out/timeline_20.csv python tools/perf_probe.py 20 600 2 1 1 2>&1 | head -8
} 2>&1 | tee gpurun_out/hog.log
