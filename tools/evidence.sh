#!/bin/bash
# Round evidence on one B200 (run through gpurun; everything lands in gpurun_out/):
#  1. clean rebuild of the library ON the GPU box + the whole -m gpu suite
#  2. ncu launch list of a short bench run (per-launch device times; cold cache, serialised: compare shares, not absolutes)
#  3. ncu --set full of the conv kernels of one step (DRAM bytes, tensor-pipe activity of the roofline kernel)
#  4. compute-sanitizer memcheck / racecheck over the smoke step and the no-allocation encoder
set -u
O=gpurun_out
mkdir -p $O
{ echo "== clean build on $(nvidia-smi --query-gpu=name --format=csv,noheader | head -1), $(date -u +%FT%TZ)"; python deep-go_b200/build.py --force; echo "build exit $?";
  timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4; } > $O/clean_build_gpu_tests.txt 2>&1
DG_NO_GRAPH=1 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active \
    --clock-control none -c 120 --csv --log-file $O/ncu_launches.csv python tools/prof_step.py 3 > $O/ncu_launches.log 2>&1
echo "ncu launches exit $?"
DG_NO_GRAPH=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_umma_kernel -s 3 -c 3 -o $O/ncu_full_conv -f python tools/prof_step.py 3 > $O/ncu_full.log 2>&1
echo "ncu full exit $?"
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python __graft_entry__.py --smoke > $O/sanitizer_memcheck_smoke.txt 2>&1
echo "memcheck smoke exit $?"
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_zz_no_alloc.py tests/test_gpu_encode.py -m gpu -q > $O/sanitizer_memcheck_encode.txt 2>&1
echo "memcheck encode exit $?"
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python __graft_entry__.py --smoke > $O/sanitizer_racecheck_smoke.txt 2>&1
echo "racecheck smoke exit $?"
tail -3 $O/clean_build_gpu_tests.txt; tail -3 $O/sanitizer_memcheck_smoke.txt; tail -3 $O/sanitizer_memcheck_encode.txt; tail -3 $O/sanitizer_racecheck_smoke.txt
