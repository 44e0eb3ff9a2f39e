#!/bin/bash
# GPT-2 medium at N GPUs under variants of the sliced step (peer route): where does the pushing step's time go?
N=${1:-8}
export DGS_PEER_TIMEOUT_S=30
mkdir -p gpurun_out
run() {
  name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29515 \
      bench.py --gpus $N --workload gpt2_medium --steps 32 --warmup 12 --no-parity --no-extra --no-microbench > gpurun_out/gpt2_$name.json 2> gpurun_out/gpt2_$name.err
  echo "$name rc=$? $(python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/gpt2_$name.json').read().strip().splitlines()[-1])
    r=d['roofline']
    print('seq/s', round(d['value'],1), 'tail_us', round(d['grad_sync_us_per_step'],1), 'step_us', round(r['us_per_launch'],1), 'launches/step', r['launches_per_step'], d['route'][:24])
except Exception as e:
    print('unparsed', e)
PY
)" | tee -a gpurun_out/r2_sweep_n${N}_gpt2.txt
}
run default
run per_bucket DGS_SLICE_LAUNCH=per_bucket
run unroll1 DGS_STEP_UNROLL=1
run split DGS_PUSH_MODE=split
run allreduce DGS_ROUTE=allreduce
