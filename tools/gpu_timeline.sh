#!/bin/bash
mkdir -p gpurun_out
FVAD_TIMELINE=gpurun_out/timeline_592.csv python tools/perf_probe.py 592 60 2 1 1 2>&1 | head -3
FVAD_TIMELINE=gpurun_out/timeline_20.csv python tools/perf_probe.py 20 240 2 1 1 2>&1 | head -3
wc -l gpurun_out/timeline_*.csv
