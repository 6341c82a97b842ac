#!/bin/bash
# final multi-GPU numbers: $1 = N
N=${1:-8}
OUT=gpurun_out/final_n$N
mkdir -p $OUT
i=0
run() {
  tag=$1; steps=$2; shift; shift
  i=$((i+1))
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29900 + i)) bench.py --gpus $N --steps $steps --warmup 5 > $OUT/$tag.json 2> $OUT/$tag.err
  echo "$tag rc=$? $(python -c "import json,sys; d=json.loads(open('$OUT/$tag.json').read().strip().splitlines()[-1]); print('dev %.4f ms  %.1f M ex/s  e2e %.4f ms  verify %s  base %s' % (d['ms_per_step'], d['value']/1e6, d['e2e']['ms_per_step'], (d.get('verify') or {}).get('ok'), (d.get('vs_torch_baseline') or {}).get('ratio')))" 2>&1 | tail -1)"
}
run async_nopf 1000 DTF_PREFETCH=0 DTF_BENCH_VERIFY=0 DTF_BENCH_BASELINE=0
run async 5000
i=$((i+1))
DTF_BENCH_BASELINE=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29900 + i)) bench.py --gpus $N --steps 1000 --warmup 5 --mode sync > $OUT/sync.json 2> $OUT/sync.err
echo "sync rc=$? $(python -c "import json,sys; d=json.loads(open('$OUT/sync.json').read().strip().splitlines()[-1]); print('dev %.4f ms  %.1f M ex/s  e2e %.4f ms  verify %s' % (d['ms_per_step'], d['value']/1e6, d['e2e']['ms_per_step'], (d.get('verify') or {}).get('ok')))" 2>&1 | tail -1)"
