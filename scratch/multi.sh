#!/bin/bash
# usage: bash scratch/multi.sh N   (on the GPU box, N GPUs): heat512 + heat256 bench lines at N ranks
N=$1
out=gpurun_out
for wl in heat512 heat256; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --workload $wl --no-cpu-baseline > $out/bench_${wl}_n$N.json 2> $out/bench_${wl}_n$N.err
  echo "$wl n$N rc=$?"; tail -c 600 $out/bench_${wl}_n$N.json
done
