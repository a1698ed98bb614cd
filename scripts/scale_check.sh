#!/bin/bash
# scaling diagnostics: per-rank step time, kernel time and evaluations per frame at N = $@ (default 2)
show() { grep "^{" | tail -1 | tee -a gpurun_out/scale_lines.jsonl | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('N',d['n_gpus'], round(d['value']), round(d['ms_per_step'],5), round(d['e2e']['value']), d['by_rank_ms_per_step_kernel_ms_evals_max_step_ms'])"; }
rm -f gpurun_out/scale_lines.jsonl
for n in ${@:-2}; do
if [ $n = 1 ]; then python bench.py --gpus 1 --steps 100 --warmup 5 --no-cpu 2>/dev/null | show; else
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 100 --warmup 5 2>/dev/null | show; fi
done
