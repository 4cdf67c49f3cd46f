#!/bin/bash
# N-GPU A/B of the graph-captured encoders under DDP (short runs)
N=${1:-2}
mkdir -p gpurun_out
for flag in "--no-graphs" ""; do
  timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
     bench.py --gpus $N --steps 10 --warmup 3 --no-sweep --no-cpu-baseline --no-eager-baseline $flag > gpurun_out/bench_n${N}g$flag.json 2> gpurun_out/bench_n${N}g$flag.err
  echo "== bench N=$N $flag exit $?"; grep -v "Warning\|warn\|^$" gpurun_out/bench_n${N}g$flag.err | tail -n 4 | cut -c1-300
  python - "gpurun_out/bench_n${N}g$flag.json" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d['value'],1), round(d['ms_per_step'],3), round(d['e2e']['value'],1), d['config'].get('encoder_cuda_graphs'), d.get('parity_check'))
    c=d.get('collectives') or {}
    print({k:v for k,v in c.items() if k!='gpu_busy' and k!='note'})
except Exception as e: print('no json', e)
P
done
