#!/bin/bash
# usage: tools/gpu_bench_variants.sh <tag> [variant ...] -- the headline bench (no CPU leg, no extras) for the main library
# and every named variant library (aslr_to_b200/variants/<variant>.so), into gpurun_out/<tag>/
tag=$1; shift
vars="$*"
/usr/local/graft/bin/gpurun --timeout 1200 -- "mkdir -p gpurun_out/$tag; for v in main $vars; do f=gpurun_out/$tag/bench_\$v.json; lib=aslr_to_b200/variants/\$v.so; if [ \$v = main ]; then lib=aslr_to_b200/libaslr_b200.so; fi; ASLR_B200_LIB=\$lib timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-extras > \$f 2>> gpurun_out/$tag/bench.err; python -c \"
import json,sys;d=json.load(open('\$f'));print('\$v',round(d['ms_per_step'],2),round(d['value']),round(d['e2e']['ms_per_step'],2),round(d['calcdiff_ms'],4),d['stages']['ms_per_solve'],d['stages']['full_batch_us'],d['stages']['floor_us'])\"; done; tail -3 gpurun_out/$tag/bench.err" 2>&1 | tail -12
