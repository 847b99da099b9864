#!/bin/bash
# Round 2, visit B: the allocation-free entry point (alone, then after every other GPU test), compute-sanitizer
# over the small parity tests, launch list + full ncu capture of the packed-add WSB kernel.
set -u
OUT=gpurun_out/r2_b
mkdir -p $OUT
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -3 $OUT/pytest_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -3 $OUT/pytest_gpu.log
PCMD="python bench.py --sentences 2000000 --steps 2 --warmup 3 --no-cpu-baseline --no-clocks"
$PCMD > $OUT/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv $PCMD > $OUT/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
$PCMD > $OUT/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:wsb_score -s 5 -c 5 -o $OUT/prof $PCMD > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -2 $OUT/ncu_full.log
for tool in memcheck racecheck synccheck; do
	timeout 1500 compute-sanitizer --tool $tool --error-exitcode 9 python -m pytest tests/test_zz_no_alloc.py tests/test_gpu_parity.py -m gpu -x -q -k "not race_free and not reference_solver" > $OUT/sanitizer_$tool.log 2>&1
	echo "sanitizer $tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" $OUT/sanitizer_$tool.log | tail -3
done
