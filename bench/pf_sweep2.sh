#!/bin/bash
N=2
OUT=gpurun_out/pf_sweep2_n$N
mkdir -p $OUT
i=0
run() {
  tag=$1; shift
  i=$((i+1))
  env "$@" DTF_BENCH_VERIFY=0 DTF_BENCH_BASELINE=0 timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29800 + i)) bench.py --gpus $N --steps 400 --warmup 5 > $OUT/$tag.json 2> $OUT/$tag.err
  echo "$tag rc=$? $(python -c "import json,sys; d=json.loads(open('$OUT/$tag.json').read().strip().splitlines()[-1]); print('dev %.4f ms  e2e %.4f ms  loss %.4f' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['last_loss']))" 2>&1 | tail -1)"
}
for k in 1 2 3; do TRIALS=3000 bash bench/tail_run.sh I$k DTF_GEMM_DBG=0; done
run off DTF_PREFETCH=0
run off2 DTF_PREFETCH=0
run b4p0 DTF_PF_BLOCKS=4 DTF_PF_POINT=0
run b6p0 DTF_PF_BLOCKS=6 DTF_PF_POINT=0
run nopair DTF_PREFETCH=0 DTF_PAIR_MASK=0
