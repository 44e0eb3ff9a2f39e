#!/bin/bash
# 2-GPU confirmation run (gpurun --gpus 2): every -m gpu test incl. the 2-slot ones, then the driver's N=2 bench command.
export DGS_PEER_TIMEOUT_S=${DGS_PEER_TIMEOUT_S:-30}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_n2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_n2_pytest.log
tail -5 gpurun_out/r2_n2_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_n2_bench.json 2> gpurun_out/r2_n2_bench.err; echo "bench rc=$?" >> gpurun_out/r2_n2_bench.err
tail -c 1500 gpurun_out/r2_n2_bench.err; head -c 600 gpurun_out/r2_n2_bench.json
