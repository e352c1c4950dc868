#!/bin/bash
# usage: tools/gpu_check.sh <tag>  -- GPU parity tests + headline bench (no CPU leg) into gpurun_out/<tag>/
tag=$1
/usr/local/graft/bin/gpurun --timeout 900 -- "mkdir -p gpurun_out/$tag; (timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/$tag/pytest.log; timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/$tag/bench.json 2> gpurun_out/$tag/bench.err; cat gpurun_out/$tag/pytest.log; python -c \"
import json;d=json.load(open('gpurun_out/$tag/bench.json'));print(d['ms_per_step'],d['value'],d['stages']['ms_per_solve'],d['stages']['full_batch_us'],d['stages']['floor_us'])\"" 2>&1 | tail -8
