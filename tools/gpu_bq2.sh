#!/bin/bash
mkdir -p gpurun_out
{
for V in "48 0" "48 1" "96 1" "120 1" "160 1" "240 1"; do
  set -- $V
  export FVAD_BQ_BLOCK=$1 FVAD_BQ_ROLL=$2
  echo "== BQ=$1 ROLL=$2"
  python -m pytest tests/test_gpu_golden.py "tests/test_gpu_denoiser.py::test_denoiser_batch_bit_exact" -x -q -m gpu 2>&1 | tail -1
  python tools/perf_probe.py 20 600 2 1 0 2>&1 | tail -1
  FVAD_SERIAL_STAGES=1 python tools/perf_probe.py 20 240 2 1 1 2>&1 | grep k_rn_biquad
done
} 2>&1 | tee gpurun_out/bq2.log
