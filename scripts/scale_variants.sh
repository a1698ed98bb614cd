#!/bin/bash
show() { grep "^{" | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['ms_per_step'],5), round(d['e2e']['value']), [[round(x[0],4), round(x[3],3)] for x in d['by_rank_ms_per_step_kernel_ms_evals_max_step_ms']])"; }
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 100 --warmup 5 2>/dev/null | show; }
echo "pin+clocks"; run
echo "nopin+clocks"; DSO_B200_NO_PIN=1 run
echo "pin+noclocks"; DSO_B200_NO_CLOCKS=1 run
echo "nopin+noclocks"; DSO_B200_NO_PIN=1 DSO_B200_NO_CLOCKS=1 run
