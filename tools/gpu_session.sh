#!/bin/bash
# One GPU-box session: parity tests, bench (both arms), ncu launch list of a short bench run.
# Usage (from the repo root, under gpurun): bash tools/gpu_session.sh [tests|bench|ncu|full ...]
mkdir -p gpurun_out
what="${*:-tests bench ncu}"
for w in $what; do
case $w in
tests)
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
  tail -25 gpurun_out/pytest_gpu.log ;;
alltests)
  timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
  tail -40 gpurun_out/pytest_gpu.log ;;
bench)
  timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref exit $?"
  cat gpurun_out/bench_reference.json
  timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
  cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err ;;
ncu)
  CMD="python bench.py --steps 5 --warmup 3 --no-sweep --no-cpu-baseline"
  timeout 600 $CMD > gpurun_out/ncu_plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_run.log 2>&1
  echo "ncu exit $?"; tail -3 gpurun_out/ncu_run.log ;;
full)
  CMD="python bench.py --steps 3 --warmup 3 --no-sweep --no-cpu-baseline"
  timeout 600 $CMD > gpurun_out/ncu_plain2.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"${NCU_KERNEL:-select_keypoints}" -s 8 -c 2 -f -o gpurun_out/prof_${NCU_KERNEL:-select_keypoints} $CMD > gpurun_out/ncu_full.log 2>&1
  echo "ncu full exit $?"; tail -3 gpurun_out/ncu_full.log ;;
esac
done
