#!/bin/bash
# one-GPU A/B sweep of executor knobs (all inside one call: boxes differ by ~4 %)
OUT=gpurun_out/sweep_n1
mkdir -p $OUT
run() {
  tag=$1; shift
  env "$@" DTF_BENCH_VERIFY=0 DTF_BENCH_BASELINE=0 timeout 100 python bench.py --gpus 1 --steps 600 --warmup 5 > $OUT/$tag.json 2> $OUT/$tag.err
  echo "$tag rc=$? $(python -c "import json,sys; d=json.loads(open('$OUT/$tag.json').read().strip().splitlines()[-1]); print('dev %.4f ms  e2e %.4f ms  mhz %s' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['clocks']['sm_mhz']))" 2>&1 | tail -1)"
}
run default
run split64 DTF_SPLIT_PUSH=64
run split48 DTF_SPLIT_PUSH=48
run pf6 DTF_PF_BLOCKS=6
run default2
