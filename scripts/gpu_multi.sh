#!/bin/bash
# One gpurun --gpus N call: the driver's own launch line for the N-GPU bench (both arms), into gpurun_out/.
N=${1:-2}
O=gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --impl reference --gpus $N --steps 20 --warmup 5 > $O/multi_ref_$N.json 2> $O/multi_ref_$N.err
echo "reference rc=$?"; cut -c1-300 $O/multi_ref_$N.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 \
  bench.py --gpus $N --steps 20 --warmup 5 > $O/multi_$N.json 2> $O/multi_$N.err
echo "b200 rc=$?"; cut -c1-400 $O/multi_$N.json; tail -5 $O/multi_$N.err
