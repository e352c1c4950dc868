#!/bin/bash
# usage: tools/gpu_check.sh <tag> [variant ...] -- GPU parity tests + headline bench (no CPU leg) into gpurun_out/<tag>/,
# then the bench alone for every named variant library (aslr_to_b200/variants/<variant>.so)
tag=$1; shift
vars="$*"
/usr/local/graft/bin/gpurun --timeout 1500 -- "mkdir -p gpurun_out/$tag; (timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -25) > gpurun_out/$tag/pytest.log; timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-extras > gpurun_out/$tag/bench.json 2> gpurun_out/$tag/bench.err; tail -6 gpurun_out/$tag/pytest.log; tail -3 gpurun_out/$tag/bench.err; for v in main $vars; do f=gpurun_out/$tag/bench.json; if [ \$v != main ]; then f=gpurun_out/$tag/bench_\$v.json; ASLR_B200_LIB=aslr_to_b200/variants/\$v.so timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-extras > \$f 2>> gpurun_out/$tag/bench.err; fi; python -c \"
import json,sys;d=json.load(open('\$f'));print('\$v',round(d['ms_per_step'],2),round(d['value']),round(d['e2e']['ms_per_step'],2),round(d['calcdiff_ms'],4),d['stages']['ms_per_solve'],d['stages']['full_batch_us'],d['stages']['floor_us'])\"; done" 2>&1 | tail -14
