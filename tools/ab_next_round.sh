#!/usr/bin/env bash
# A/B runs for the switches that were written without GPU time left.  Usage on a GPU box:
#   bash tools/ab_next_round.sh 1        # 1 GPU: native host loop on/off
#   bash tools/ab_next_round.sh 8        # 8 GPUs: native host loop and the single-bucket switch
# Prints one line per arm: the bench's value, ms/step and grad-sync tail.  Compare arms of the same invocation only.
set -u
N=${1:-1}
run() {  # label, env assignments...
  local label=$1; shift
  local launch=(python bench.py)
  if [ "$N" -gt 1 ]; then
    launch=(python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29517 bench.py)
  fi
  env "$@" "${launch[@]}" --gpus "$N" --steps 60 --warmup 10 2>/dev/null | python -c "
import json, sys
for line in sys.stdin:
    line = line.strip()
    if line.startswith('{'):
        d = json.loads(line)
        print('$label', round(d['value'], 1), d['unit'], round(d['ms_per_step'], 3), 'ms/step', 'e2e', round(d['e2e']['value'], 1), d.get('grad_sync', ''))
"
}
python -c "from determined_clone_b200 import _build; _build.build_host()"
for rep in 1 2; do
  run "python-loops  " DGS_HOST_EXT=0
  run "native-loop   " DGS_HOST_EXT=1
  if [ "$N" -gt 1 ]; then
    run "one-bucket    " DGS_HOST_EXT=0 DGS_ONE_BUCKET_BELOW_MB=64
    run "both          " DGS_HOST_EXT=1 DGS_ONE_BUCKET_BELOW_MB=64
  fi
done
