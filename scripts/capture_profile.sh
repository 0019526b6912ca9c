set -e
python bench.py --steps 2 --warmup 3 > gpurun_out/plain_r1f.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_r1f.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu1_r1f.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'pyramid_kernel|fast_kernel|quadtree_kernel|describe_kernel' -s 12 -c 4 -f -o gpurun_out/prof_r1f python bench.py --steps 2 --warmup 3 > gpurun_out/ncu2_r1f.log 2>&1
tail -c 600 gpurun_out/plain_r1f.log
