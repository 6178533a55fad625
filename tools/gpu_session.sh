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
  CMD="python bench.py --steps 6 --warmup 3 --no-sweep --no-cpu-baseline"
  timeout 600 $CMD > gpurun_out/ncu_plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_run.log 2>&1
  echo "ncu exit $?"; tail -3 gpurun_out/ncu_run.log ;;
full)
  CMD="python bench.py --steps 6 --warmup 3 --no-sweep --no-cpu-baseline"
  timeout 600 $CMD > gpurun_out/ncu_plain2.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_plain2.log; continue; }
  for K in ${NCU_KERNELS:-harris_nms describe_kernel match_to_map_slots select_keypoints}; do
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -f -o gpurun_out/prof_$K $CMD > gpurun_out/ncu_full_$K.log 2>&1
    echo "ncu full $K exit $?"; tail -2 gpurun_out/ncu_full_$K.log
  done ;;
esac
done
