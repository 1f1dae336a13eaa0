#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_denoiser.py tests/test_gpu_golden.py -x -q -m gpu > gpurun_out/quick_test.log 2>&1; echo "pytest rc=$?" >> gpurun_out/quick_test.log
tail -5 gpurun_out/quick_test.log
{
for S in 148 296 592 1184; do python tools/perf_probe.py $S 120 2 1 0 2>&1 | tail -1; done
echo "== lanes off"; FVAD_BIQUAD_LANES=0 python tools/perf_probe.py 592 120 2 1 0 2>&1 | tail -1
FVAD_SERIAL_STAGES=1 python tools/perf_probe.py 592 120 2 1 1 2>&1 | tail -26 | head -12
} 2>&1 | tee gpurun_out/quick_perf.log
