#!/bin/bash
# Run on the GPU box (under gpurun): launch list + one full ncu capture of the top kernel and of the STFT kernel.
# Every ncu command is preceded by the same command without ncu (must exit 0).
set -x
mkdir -p gpurun_out
SHORT="python bench.py --seconds 600 --steps 1 --warmup 1 --no-e2e --no-cpu --no-batched --no-strong --no-sweep"
$SHORT > gpurun_out/short_plain.json 2> gpurun_out/short_plain.err || { echo "plain run failed"; tail -5 gpurun_out/short_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $SHORT > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_rn_biquad_chain -s 1 -c 1 -o gpurun_out/prof_biquad $SHORT > gpurun_out/ncu_biquad.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_stft2048 -s 12 -c 1 -o gpurun_out/prof_stft $SHORT > gpurun_out/ncu_stft.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_rn_rnn -s 1 -c 1 -o gpurun_out/prof_rnn $SHORT > gpurun_out/ncu_rnn.log 2>&1
# the bandwidth-bound STFT launch on its own (configs[2] chunk: 4096 streams x 1406 frames)
python tools/stft_probe.py > gpurun_out/stft_probe.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stft2048 -c 1 -o gpurun_out/prof_stft_4096x1406 python tools/stft_probe.py > gpurun_out/ncu_stft_big.log 2>&1
ls -la gpurun_out
