#!/bin/bash
# GPU box: parity of every RNN kernel shape, then throughput at 296 / 592 / 1184 streams per shape.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_denoiser.py -x -q -m gpu > gpurun_out/rnn_variants_test.log 2>&1; echo "pytest rc=$?" >> gpurun_out/rnn_variants_test.log
tail -3 gpurun_out/rnn_variants_test.log
for S in 296 592; do
  for RG in "default default" "2 1" "4 1" "1 2" "1 3" "2 2" "2 3" "4 2" "4 3"; do
    set -- $RG
    if [ "$1" = default ]; then unset FVAD_RNN_R FVAD_RNN_G; else export FVAD_RNN_R=$1 FVAD_RNN_G=$2; fi
    echo "== S=$S R=$1 G=$2"
    python tools/perf_probe.py $S 120 2 1 0 2>&1 | tail -1
  done
done 2>&1 | tee gpurun_out/rnn_variants.log
