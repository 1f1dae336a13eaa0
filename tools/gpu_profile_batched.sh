#!/bin/bash
# Run on the GPU box (under gpurun): throughput regime (592 streams) -- plain timing, serialised per-kernel shares,
# then one full ncu capture each of the RNN (R=4), pitch, synthesis and pitch-feature kernels.
set -x
mkdir -p gpurun_out
python tools/perf_probe.py 592 120 2 1 0 > gpurun_out/p592.log 2>&1 || { tail -5 gpurun_out/p592.log; exit 1; }
FVADK_SERIAL_STAGES=1 python tools/perf_probe.py 592 120 2 1 1 > gpurun_out/p592_serial.log 2>&1
for k in k_rn_rnn k_rn_pitch k_rn_synth k_rn_pfeat k_rn_analysis k_rn_biquad; do
  timeout 600 ncu --set full --clock-control none --import-source on -k "regex:^${k}(<.*)?\$" -s 1 -c 1 -f -o gpurun_out/t_$k python tools/perf_probe.py 592 60 1 > gpurun_out/ncu_t_$k.log 2>&1
done
cat gpurun_out/p592.log gpurun_out/p592_serial.log
ls -la gpurun_out
