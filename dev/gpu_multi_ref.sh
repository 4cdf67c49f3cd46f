#!/bin/bash
# the reference arm under torchrun (rank 0 alone runs it): gpurun --gpus 2 -- 'bash dev/gpu_multi_ref.sh 2'
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 \
   bench.py --impl reference --gpus $N --steps 2 --warmup 1 --cpu-budget 40 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err
echo "== ref N=$N exit $?"; tail -c 1500 gpurun_out/bench_ref_n$N.json; grep -v "^W1\|^\*\|OMP" gpurun_out/bench_ref_n$N.err | tail -n 5
