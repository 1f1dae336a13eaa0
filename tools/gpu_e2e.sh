#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_fullsize.py 2>&1 | tail -2
python bench.py --seconds 1200 --steps 2 --warmup 1 --no-cpu --no-batched 2>gpurun_out/e2e.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('value',d['value'],d['ms_per_step']); print('e2e',d['e2e']['value'],d['e2e']['ms_per_step'],d['e2e']['segments_equal_device_resident_run'])"
