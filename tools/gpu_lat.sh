#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_denoiser.py tests/test_gpu_golden.py -x -q -m gpu > gpurun_out/quick_test.log 2>&1; echo "pytest rc=$?" >> gpurun_out/quick_test.log
tail -3 gpurun_out/quick_test.log
{
for S in 20 40 80; do
  for P in default "0,5,4,3,2,1" "0,3,2,1,1,0"; do
    echo "== S=$S FVAD_PRIO=$P"; if [ "$P" = default ]; then unset FVAD_PRIO; else export FVAD_PRIO=$P; fi
    python tools/perf_probe.py $S 600 2 1 0 2>&1 | tail -1
  done
done
unset FVAD_PRIO
FVAD_SERIAL_STAGES=1 python tools/perf_probe.py 20 240 2 1 1 2>&1 | tail -26 | head -6
} 2>&1 | tee gpurun_out/lat.log
