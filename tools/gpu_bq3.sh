#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_denoiser.py tests/test_gpu_golden.py -x -q -m gpu 2>&1 | tail -2
{
python tools/perf_probe.py 20 600 2 1 0 2>&1 | tail -1
FVAD_SERIAL_STAGES=1 python tools/perf_probe.py 20 240 2 1 1 2>&1 | grep -E "k_rn_biquad|rep 1"
for S in 80 592; do python tools/perf_probe.py $S 120 2 1 0 2>&1 | tail -1; done
} 2>&1 | tee gpurun_out/bq3.log
