#!/bin/bash
OUT=gpurun_out/sweep_n1
mkdir -p $OUT
run() {
  tag=$1; shift
  env "$@" DTF_BENCH_VERIFY=0 DTF_BENCH_BASELINE=0 timeout 120 python bench.py --gpus 1 --steps 1000 --warmup 5 > $OUT/$tag.json 2> $OUT/$tag.err
  echo "$tag rc=$? $(python -c "import json,sys; d=json.loads(open('$OUT/$tag.json').read().strip().splitlines()[-1]); print('dev %.4f ms  e2e %.4f ms  loss %.4f' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['last_loss']))" 2>&1 | tail -1)"
}
run default
run nopf DTF_PREFETCH=0
run split56 DTF_SPLIT_PUSH=56
run split56nopf DTF_SPLIT_PUSH=56 DTF_PREFETCH=0
run split40 DTF_SPLIT_PUSH=40
run pf2 DTF_PF_BLOCKS=2
run default2
