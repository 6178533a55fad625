#!/bin/bash
# ncu evidence of one round (run under gpurun from the repo root): launch list of a short bench run + one
# `--set full` capture per kernel. Each ncu pass only after the same command exited 0 without ncu.
# The headline config runs first: 5 keyframe launches of 2 images, then 64-image batches; -s 8 skips into those.
mkdir -p gpurun_out
rm -f gpurun_out/*.ncu-rep gpurun_out/launches.csv
CMD="python bench.py --steps 6 --warmup 3 --no-cpu-baseline --only euroc_sequence"
timeout 600 $CMD > gpurun_out/ncu_plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/ncu_plain.log; exit 1; }
if [ -z "$NCU_NO_LIST" ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_run.log 2>&1; echo "launch list exit $?"
fi
for K in ${NCU_KERNELS-harris_nms_kernel order_candidates_kernel select_keypoints_wave_kernel subpixel_kernel describe_kernel back_project_slots_kernel project_landmarks_slots_kernel prune_observations_slots_kernel pack_pool_slots_kernel bin_projections_kernel match_to_map_binned_kernel match_stats_kernel uninit_filter_kernel match_uninitialised_kernel hamming_best_lb}; do
  S=8; [ "$K" = hamming_best_lb ] && S=2   # (the sweep's first launches are the i.i.d. input of SURVEY 8(d))
  timeout 900 ncu --set full --clock-control none --import-source on --warp-sampling-interval 0 -k regex:$K -s $S -c 1 -f -o gpurun_out/prof_$K $CMD > gpurun_out/ncu_full_$K.log 2>&1; echo "ncu full $K exit $?"
done
# the 2D-2D matchers run in other configs (the headline config runs first there too and has none of these kernels)
if [ -z "$NCU_NO_2D" ]; then
  CMD2="python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-sweep --only stereo_pair"
  CMD3="python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-sweep --only d455_sequence"
  timeout 600 $CMD2 > gpurun_out/ncu_plain2.log 2>&1 && for K in stereo_filter_kernel match_stereo_kernel; do
    timeout 900 ncu --set full --clock-control none --import-source on --warp-sampling-interval 0 -k regex:$K -s 4 -c 1 -f -o gpurun_out/prof_${K}_stereo_pair $CMD2 > gpurun_out/ncu_full_${K}_sp.log 2>&1; echo "ncu full $K (stereo pair) exit $?"
  done
  timeout 600 $CMD3 > gpurun_out/ncu_plain3.log 2>&1 && for K in stereo_filter_kernel match_stereo_kernel; do
    timeout 900 ncu --set full --clock-control none --import-source on --warp-sampling-interval 0 -k regex:$K -s 4 -c 1 -f -o gpurun_out/prof_${K}_motion_d455 $CMD3 > gpurun_out/ncu_full_${K}_d455.log 2>&1; echo "ncu full $K (d455 motion) exit $?"
  done
fi
ls -la gpurun_out/*.ncu-rep
