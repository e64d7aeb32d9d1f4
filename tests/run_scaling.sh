#!/bin/bash
# Scaling runs kept under profiles/ (round 2): bash tests/run_scaling.sh "<configs>" "<N list>" [extra bench flags]
#   e.g. gpurun --gpus 8 -- 'bash tests/run_scaling.sh "headline c3 c4" "4 8"'
# One JSON line per (config, N) in gpurun_out/r02_scale_<config>_n<N><tag>.json; stderr beside it.
CONFIGS=${1:-headline}
NS=${2:-1}
shift 2
TAG=${SCALE_TAG:-}
PORT=29510
for cfg in $CONFIGS; do
  for n in $NS; do
    out=gpurun_out/r02_scale_${cfg}_n${n}${TAG}
    PORT=$((PORT + 1))
    if [ "$n" = "1" ]; then
      timeout 420 python bench.py --gpus 1 --config $cfg --no-legs --no-cpu-baseline "$@" > $out.json 2> $out.err
    else
      timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT \
        bench.py --gpus $n --config $cfg --no-legs --no-cpu-baseline "$@" > $out.json 2> $out.err
    fi
    echo "$cfg N=$n rc=$? $(cut -c1-160 $out.json)"
  done
done
