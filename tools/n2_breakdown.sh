run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 --steps 50 --warmup 5 --no-microbench 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('$2', round(d['ms_per_step'],3), 'ms/step', round(d['value']), 'samples/s', 'tail_us', d.get('grad_sync_us_per_step'))"; }
run 29601 full
DGS_DEBUG_SKIP=buffers run 29602 no_buffer_bcast
DGS_DEBUG_SKIP=allreduce run 29603 no_allreduce
DGS_DEBUG_SKIP=buffers,allreduce run 29604 neither
NCCL_NTHREADS=64 run 29605 full_again
