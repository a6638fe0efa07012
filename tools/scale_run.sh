#!/bin/bash
# bench.py at the GPU counts given on the command line (default 1 2 4 8), launched the way the driver launches it (one rank
# per GPU, torchrun).  BASELINE configs[2]: ONE 10 M-segment input sharded over the ranks (strong scaling).
set -u
mkdir -p gpurun_out
for n in ${@:-1 2 4 8}; do
  if [ $n -eq 1 ]; then python bench.py --gpus 1 --no-cpu --no-secondary > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err; fi
  python - <<PY
import json
d = json.loads([l for l in open('gpurun_out/scale_$n.json') if l.startswith('{')][-1])
w = d.get('weak_scaling') or {}
print($n, 'value', round(d['value']/1e6, 1), 'M/s  e2e', round(d['e2e']['value']/1e6, 1), 'M/s  pageable',
      round(d['e2e']['pageable']['value']/1e6, 1), ' ms', round(d['ms_per_step'], 3), d['clocks']['sm_mhz'], d['clocks']['reasons'],
      d['check']['sha256'][:10], d['check']['sharded_equals_single_gpu'], '| weak value', round(w.get('value', 0)/1e6, 1),
      'e2e', round((w.get('e2e') or {}).get('value', 0)/1e6, 1), w.get('every_rank_equals_the_single_gpu_result'))
PY
done
nproc
