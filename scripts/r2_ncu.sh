#!/usr/bin/env bash
# round-2 ncu evidence: launch list of a short bench run and one --set full capture of the step's kernels (one steady-state step)
export NGPB_BENCH_PRETRAIN=40 NGPB_BENCH_NO_BASELINE=1
CMD="python bench.py --steps 2 --warmup 3"
mkdir -p gpurun_out
$CMD > gpurun_out/plain.log 2> gpurun_out/plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 400 --csv --log-file gpurun_out/r02_nerf_launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2> gpurun_out/plain2.err && \
ncu --set full --clock-control none --import-source on -k regex:'fused_net_kernel|adam_ema_step_kernel|grid_backward_kernel|grid_forward_kernel|fused_grid_net_kernel|generate_training_samples_kernel|compute_loss_kernel|gather_tiles_kernel' -s 368 -c 9 -o gpurun_out/r02_nerf_step $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu1.log gpurun_out/ncu2.log; ls -la gpurun_out/*.ncu-rep gpurun_out/r02_nerf_launches.csv
