#!/bin/bash
# Round-end style validation on the GPU box: full GPU test suite, both bench arms, ncu launch list + captures (latency
# regime = the bench's own launches; throughput regime = 592 streams).
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/final_test.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final_test.log
tail -3 gpurun_out/final_test.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; tail -1 gpurun_out/final_smoke.log
python bench.py --impl reference > gpurun_out/bench_reference_n1.json 2> gpurun_out/bench_reference.err; tail -c 600 gpurun_out/bench_reference_n1.json
python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench.err; tail -c 3000 gpurun_out/bench_n1.json
bash tools/gpu_profile.sh > gpurun_out/profile.log 2>&1
bash tools/gpu_profile_batched.sh > gpurun_out/profile_batched.log 2>&1
tail -8 gpurun_out/p592_serial.log
ls gpurun_out
