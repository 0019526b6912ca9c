#!/bin/bash
# compute-sanitizer over the parity tests (race detection / memory checking for every kernel).
#   bash scripts/sanitize.sh            on the GPU box; summaries land in gpurun_out/sanitize_*.log, the digest in gpurun_out/sanitize_summary.txt
# memcheck: out-of-bounds / misaligned accesses in every kernel the selected tests launch (extraction at the three BASELINE
# resolutions is too slow under the tool, so the VGA config, the small golden, the batch path, the Hamming kernels, the frame grid,
# the device matcher and the no-allocation entry point are run).  racecheck: shared-memory hazards — the describe kernel overlays
# its horizontal blur pass on the raw window under __syncwarp, the quadtree sort ping-pongs halves of one shared buffer, the FAST
# kernel reuses the image tile for its survivor lists: exactly what the tool exists for.
# NOTE: on the GPU pool this repository is developed on, compute-sanitizer is closed (it prints a notice and exits without running
# anything); the script is kept for machines where the tool is available.  What stands in for it here: tests/test_gpu_stress.py
# (hazard-prone kernels repeated under changing co-residency must give bit-identical results) and the bit-exact parity suite.
set -u
OUT=gpurun_out
mkdir -p $OUT
SEL_MEM='tests/test_gpu_extract.py::test_against_committed_reference_golden tests/test_gpu_extract.py::test_batch_equals_single_and_is_deterministic tests/test_gpu_hamming.py tests/test_gpu_framegrid.py tests/test_gpu_matcher_dev.py::test_whole_function_equals_reference_goldens tests/test_gpu_mappoint.py tests/test_zz_no_alloc.py'
SEL_RACE='tests/test_gpu_extract.py::test_against_committed_reference_golden tests/test_gpu_hamming.py::test_knn2_matches_oracle tests/test_gpu_matcher_dev.py::test_whole_function_equals_reference_goldens tests/test_zz_no_alloc.py::test_noalloc_equals_existing_entry_point_twice_in_a_row'
: > $OUT/sanitize_summary.txt
for tool in memcheck racecheck; do
    if [ $tool = memcheck ]; then SEL=$SEL_MEM; else SEL=$SEL_RACE; fi
    timeout 1500 compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 99 python -m pytest $SEL -m gpu -x -q > $OUT/sanitize_$tool.log 2>&1
    rc=$?
    echo "== $tool: exit code $rc" >> $OUT/sanitize_summary.txt
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|error" $OUT/sanitize_$tool.log | tail -6 >> $OUT/sanitize_summary.txt
done
cat $OUT/sanitize_summary.txt
