#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_denoiser.py -x -q -m gpu > gpurun_out/quick_test.log 2>&1; echo "pytest rc=$?" >> gpurun_out/quick_test.log
tail -3 gpurun_out/quick_test.log
{
for BQ in 32 48 60 80; do
  echo "== FVAD_BQ_BLOCK=$BQ"; FVAD_BQ_BLOCK=$BQ python tools/perf_probe.py 20 600 2 1 0 2>&1 | tail -1
  FVAD_BQ_BLOCK=$BQ FVAD_SERIAL_STAGES=1 python tools/perf_probe.py 20 240 2 1 1 2>&1 | grep k_rn_biquad
done
} 2>&1 | tee gpurun_out/bq.log
