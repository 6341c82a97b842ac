#!/bin/bash
# prefetch on/off comparison at N GPUs (N = $1, default 1); results under gpurun_out/pf_<tag>/
N=${1:-1}
OUT=gpurun_out/pf_n$N
mkdir -p $OUT
run() {  # tag, env...
  tag=$1; shift
  if [ "$N" = 1 ]; then
    env "$@" DTF_BENCH_VERIFY=0 DTF_BENCH_BASELINE=0 timeout 300 python bench.py --gpus 1 --steps 400 --warmup 5 > $OUT/$tag.json 2> $OUT/$tag.err
  else
    env "$@" DTF_BENCH_VERIFY=0 DTF_BENCH_BASELINE=0 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 400 --warmup 5 > $OUT/$tag.json 2> $OUT/$tag.err
  fi
  echo "$tag rc=$? $(python -c "import json,sys; d=json.loads(open('$OUT/$tag.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['e2e']['last_loss'])" 2>&1 | tail -1)"
}
run off DTF_PREFETCH=0
run pf2 DTF_PF_BLOCKS=2
run pf1 DTF_PF_BLOCKS=1
run pf4 DTF_PF_BLOCKS=4
