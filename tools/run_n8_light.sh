#!/bin/bash
N=${1:-8}
tr() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}" 2>gpurun_out/n${N}_err.log; }
echo "== bench cifar ours (micro 1024 MiB)"; tr 29802 bench.py --gpus $N --steps 50 --warmup 8 --micro-mb 1024 | tee gpurun_out/bench_r1_n${N}_v2.json | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(round(d['value']), d['unit'], round(d['ms_per_step'],3), 'ms/step e2e', round(d['e2e']['value']), 'tail_us', d['grad_sync_us_per_step'], 'micro', d['gradsync_microbench'])"
echo "== micro 256"; tr 29803 bench.py --gpus $N --steps 6 --warmup 3 --micro-mb 256 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d['gradsync_microbench'])"
echo "== gpt2 ours"; tr 29850 bench.py --workload gpt2_medium --gpus $N --steps 24 --warmup 8 --no-microbench | tee gpurun_out/bench_r1_gpt2_n${N}_v2.json | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(round(d['value'],1), d['unit'], round(d['ms_per_step'],2), 'ms/step tail_us', d['grad_sync_us_per_step'], 'roofline', d['roofline'] and round(d['roofline']['frac'],3))"
