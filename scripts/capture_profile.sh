#!/bin/bash
# ncu evidence of the bench's headline step (B200_PROFILING.md recipe).  Usage on the GPU box:
#   bash scripts/capture_profile.sh <tag> [extra bench flags]
# 1. plain run (must exit 0), 2. launch list (gpu__time_duration only), 3. one `--set full` capture of the four kernels of a
# warm step (ten launches at the bench's batch sizes: seven pyramid levels, FAST, quadtree, describe), exported as the raw CSV that
# profiles/ keeps.  Numbers printed by the runs under ncu are never bench values.
set -e
TAG=$1; shift || true
OUT=gpurun_out
mkdir -p $OUT
python bench.py --quick --steps 2 --warmup 3 "$@" > $OUT/plain_$TAG.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --quick --steps 2 --warmup 3 "$@" > $OUT/ncu1_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'pyramid_kernel|pyramid_level_kernel|fast_kernel|quadtree_kernel|describe_kernel' -s 30 -c 10 -f \
    -o $OUT/prof_$TAG python bench.py --quick --steps 2 --warmup 3 "$@" > $OUT/ncu2_$TAG.log 2>&1
ncu -i $OUT/prof_$TAG.ncu-rep --page raw --csv > $OUT/${TAG}_ncu_full_raw.csv 2>/dev/null
tail -c 400 $OUT/plain_$TAG.log
