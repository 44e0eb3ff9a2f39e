#!/bin/bash
# Parameter-exchange / push-kernel variants of the peer route on N GPUs, 1 GiB fp32 AdamW + clip microbench.
N=${1:-8}
export DGS_PEER_TIMEOUT_S=30
mkdir -p gpurun_out
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 \
      bench.py --gpus $N --micro-only --micro-mb 1024 --micro-dtypes f32 --micro-aggs 1 > gpurun_out/sweep_$name.json 2> gpurun_out/sweep_$name.err
  echo "$name rc=$? $(python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/sweep_$name.json').read().strip().splitlines()[-1])
    r=d['gradsync_microbench']['rows'][0]
    print(round(r['us_per_step'],1), round(r['frac'],3), r['phases_us'])
except Exception as e:
    print('unparsed', e)
PY
)" | tee -a gpurun_out/r2_sweep_n${N}_push.txt
}
run fused_rot DGS_PUSH_MODE=fused
run split DGS_PUSH_MODE=split
run fused_unroll2 DGS_PUSH_MODE=fused DGS_STEP_UNROLL=2
run split_flat4 DGS_PUSH_MODE=split DGS_FLAT_UNROLL=4
run split_flat1 DGS_PUSH_MODE=split DGS_FLAT_UNROLL=1
