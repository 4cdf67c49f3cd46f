#!/bin/bash
# N-GPU bench of the final state (graph-captured encoders; parity legs on; no sweeps / baselines)
N=${1:-8}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
   bench.py --gpus $N --steps 10 --warmup 3 --no-sweep --no-cpu-baseline --no-eager-baseline > gpurun_out/bench_n${N}.json 2> gpurun_out/bench_n${N}.err
echo "== bench N=$N exit $?"; grep -v "Warning\|warn\|^$\|\*\*\*\|OMP_NUM" gpurun_out/bench_n${N}.err | tail -n 4 | cut -c1-300
python - "gpurun_out/bench_n${N}.json" <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(round(d['value'],1), round(d['ms_per_step'],3), round(d['e2e']['value'],1), d['config'].get('encoder_cuda_graphs'))
print(d.get('parity_check'))
c=d.get('collectives') or {}
print({k:v for k,v in c.items() if k!='note'})
P
