#!/bin/bash
# N independent host-to-host runs side by side, one per GPU (what N ranks of bench.py do), for chunk-size experiments:
#   N=4 OS1_CHUNK=8 bash scripts/e2e_multi.sh
N=${N:-4}
for i in $(seq 0 $((N-1))); do
  CUDA_VISIBLE_DEVICES=$i RES=1080p B=256 STEPS=20 python scripts/e2e.py > /tmp/e2e_$i.log 2>&1 &
done
wait
for i in $(seq 0 $((N-1))); do grep "^e2e" /tmp/e2e_$i.log; done
