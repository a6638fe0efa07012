#!/bin/bash
# 1/2/4/8-GPU weak-scaling run of bench.py, launched the way the driver launches it (one rank per GPU, torchrun).
for n in 1 2 4 8; do
  if [ $n -eq 1 ]; then python bench.py --gpus 1 --no-cpu > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err; fi
  python -c "
import json,sys
d=json.loads([l for l in open('gpurun_out/scale_$n.json') if l.startswith('{')][-1])
print($n, 'value', round(d['value']/1e6,1), 'M/s  e2e', round(d['e2e']['value']/1e6,1), 'M/s  ms', round(d['ms_per_step'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
nproc
