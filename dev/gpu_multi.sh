#!/bin/bash
# N-GPU checks: NCCL parity test, then the DDP bench.  gpurun --gpus N --timeout 900 -- 'bash dev/gpu_multi.sh N'
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/gpus_$N.txt
timeout 400 python -m pytest tests/test_dist_gpu.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/test_dist_gpu.log 2>&1; rc=$?
echo "== dist test exit $rc"; tail -n 15 gpurun_out/test_dist_gpu.log
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
   bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "== bench N=$N exit $?"; tail -c 2500 gpurun_out/bench_n$N.json; tail -n 5 gpurun_out/bench_n$N.err
exit $rc
