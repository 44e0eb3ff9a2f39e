#!/bin/bash
# One 8-GPU session: NCCL probe, scaling bench (ours vs torch DDP), grad-sync microbench by size and bucket size.
N=${1:-8}
tr() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}" 2>gpurun_out/n${N}_err.log; }
micro() { python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); m=d['gradsync_microbench']; print('$1', 'N=$N', m['bucket_mib'], 'MiB', round(m['us_per_step'],1), 'us', round(m['frac_of_roofline'],3), m['bound'], round(m['roofline_us'],1))"; }
echo "== nccl probe"; tr 29801 tools/nccl_probe.py | tee gpurun_out/nccl_probe_n${N}.jsonl
echo "== bench cifar ours"; tr 29802 bench.py --gpus $N --steps 50 --warmup 8 --micro-mb 256 | tee gpurun_out/bench_r1_n${N}.json | cut -c1-250; tail -2 gpurun_out/n${N}_err.log
echo "== bench cifar torch"; tr 29803 bench.py --impl torch --gpus $N --steps 50 --warmup 8 | tail -1
for b in 25 64 128; do for mb in 256 1024 2048; do DGS_BUCKET_MB=$b tr 298$((b%10))$((mb%9)) bench.py --gpus $N --steps 6 --warmup 3 --micro-mb $mb | micro "bucket=${b}MiB"; done; done
echo "== gpt2 ours"; tr 29850 bench.py --workload gpt2_medium --gpus $N --steps 24 --warmup 8 --no-microbench | tee gpurun_out/bench_r1_gpt2_n${N}.json | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(round(d['value'],1), d['unit'], round(d['ms_per_step'],2), 'ms/step tail_us', d['grad_sync_us_per_step'], 'roofline', d['roofline'] and round(d['roofline']['frac'],3))"
echo "== gpt2 torch"; tr 29851 bench.py --impl torch --workload gpt2_medium --gpus $N --steps 24 --warmup 8 | tail -1
