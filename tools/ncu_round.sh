#!/bin/bash
# ncu evidence of one round (run under gpurun from the repo root): launch list of a short bench run + one
# `--set full` capture per kernel. Each ncu pass only after the same command exited 0 without ncu.
mkdir -p gpurun_out
rm -f gpurun_out/*.ncu-rep
CMD="python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/ncu_plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/ncu_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_run.log 2>&1; echo "launch list exit $?"
for K in ${NCU_KERNELS:-harris_nms_kernel order_candidates_kernel select_keypoints_kernel describe_kernel bin_projections_kernel match_to_map_binned_kernel match_stats_kernel hamming_best_lb}; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -f -o gpurun_out/prof_$K $CMD > gpurun_out/ncu_full_$K.log 2>&1; echo "ncu full $K exit $?"
done
ls -la gpurun_out/*.ncu-rep
