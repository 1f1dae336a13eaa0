#!/bin/bash
# Both bench arms, then the ncu launch list + captures (latency regime = the bench's own launches; throughput regime = 592
# streams), summarised ON the box: the .ncu-rep files (with sources) exceed what gpurun copies back.
mkdir -p gpurun_out
python bench.py --impl reference > gpurun_out/bench_reference_n1.json 2> gpurun_out/bench_reference.err
python bench.py --batched-streams 1184,148,296,592,2368 > gpurun_out/bench_n1.json 2> gpurun_out/bench.err; tail -c 1500 gpurun_out/bench_n1.json
bash tools/gpu_profile.sh > gpurun_out/profile.log 2>&1
bash tools/gpu_profile_batched.sh > gpurun_out/profile_batched.log 2>&1
python tools/ncu_summarise.py ${TAG:-r02f} gpurun_out > gpurun_out/summarise.log 2>&1
cp profiles/${TAG:-r02f}_* profiles/ncu_traffic.json gpurun_out/
rm -f gpurun_out/*.ncu-rep
du -sh gpurun_out; ls gpurun_out
