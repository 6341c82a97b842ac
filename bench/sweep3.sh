#!/bin/bash
N=${1:-2}
OUT=gpurun_out/sweep3_n$N
mkdir -p $OUT
i=0
run() {
  tag=$1; shift
  i=$((i+1))
  if [ "$N" = 1 ]; then
    env "$@" DTF_BENCH_VERIFY=0 DTF_BENCH_BASELINE=0 timeout 120 python bench.py --gpus 1 --steps 1000 --warmup 5 > $OUT/$tag.json 2> $OUT/$tag.err
  else
    env "$@" DTF_BENCH_VERIFY=0 DTF_BENCH_BASELINE=0 timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29800 + i)) bench.py --gpus $N --steps 1000 --warmup 5 > $OUT/$tag.json 2> $OUT/$tag.err
  fi
  echo "$tag rc=$? $(python -c "import json,sys; d=json.loads(open('$OUT/$tag.json').read().strip().splitlines()[-1]); print('dev %.4f ms  e2e %.4f ms  loss %.4f' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['last_loss']))" 2>&1 | tail -1)"
}
run base DTF_LATE_WGRAD=0
run late DTF_LATE_WGRAD=1
run late48 DTF_LATE_WGRAD=1 DTF_SPLIT_PUSH=48
run late64 DTF_LATE_WGRAD=1 DTF_SPLIT_PUSH=64
run late40 DTF_LATE_WGRAD=1 DTF_SPLIT_PUSH=40
run base2 DTF_LATE_WGRAD=0
