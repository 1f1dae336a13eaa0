#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_fullsize.py > gpurun_out/quick_test.log 2>&1; echo "pytest rc=$?" >> gpurun_out/quick_test.log
tail -3 gpurun_out/quick_test.log
{
for F in 4 6 8; do echo "== FVAD_PITCH_F=$F"; FVAD_PITCH_F=$F python tools/perf_probe.py 592 120 2 1 0 2>&1 | tail -1; done
for S in 20 80 148 296 1184; do python tools/perf_probe.py $S 120 2 1 0 2>&1 | tail -1; done
FVAD_SERIAL_STAGES=1 python tools/perf_probe.py 592 120 2 1 1 2>&1 | tail -26
} 2>&1 | tee gpurun_out/quick_perf.log
