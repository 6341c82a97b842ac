#!/bin/bash
OUT=gpurun_out/sweep_n1c
mkdir -p $OUT
run() {
  tag=$1; shift
  env "$@" DTF_BENCH_VERIFY=0 DTF_BENCH_BASELINE=0 timeout 120 python bench.py --gpus 1 --steps 1000 --warmup 5 > $OUT/$tag.json 2> $OUT/$tag.err
  echo "$tag rc=$? $(python -c "import json,sys; d=json.loads(open('$OUT/$tag.json').read().strip().splitlines()[-1]); print('dev %.4f ms  e2e %.4f ms  mhz %s' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['clocks']['sm_mhz']))" 2>&1 | tail -1)"
}
run base
run stream DTF_APPLY_STREAM=1
run base2
run stream2 DTF_APPLY_STREAM=1
