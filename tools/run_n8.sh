#!/bin/bash
# N-GPU confirmation run (gpurun --gpus N): the driver's bench command under torchrun.
N=${1:-8}
export DGS_PEER_TIMEOUT_S=${DGS_PEER_TIMEOUT_S:-30}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_n${N}_bench.json 2> gpurun_out/r2_n${N}_bench.err; echo "bench rc=$?" >> gpurun_out/r2_n${N}_bench.err
tail -c 1200 gpurun_out/r2_n${N}_bench.err; head -c 400 gpurun_out/r2_n${N}_bench.json
