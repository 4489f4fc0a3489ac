#!/bin/bash
# ncu evidence for the NeRF training step (run on the GPU box through gpurun):
#   1) the bench command exits 0 on its own, 2) launch list (durations only), 3) one --set full capture
#   of a steady-state step, exported as CSV pages. Numbers printed under ncu are never bench values.
set -u
export NGPB_BENCH_PRETRAIN=300
K='regex:generate_training|emit_training|grid_forward|grid_backward|fused_net|reduce_dw|compute_loss|fill_rollover|adam_ema|reduce_loss|bitfield|grid_samples|mean_density|mark_untrained'
OUT=gpurun_out
mkdir -p $OUT
python bench.py --steps 4 --warmup 3 > $OUT/ncu_plain.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/ncu_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" --launch-skip 3000 --launch-count 700 --csv --log-file $OUT/r01_nerf_launches.csv \
    python bench.py --steps 4 --warmup 3 > $OUT/ncu_list.log 2>&1
echo "launch list rows: $(wc -l < $OUT/r01_nerf_launches.csv)"
ncu --set full --import-source on --clock-control none -k "$K" --launch-skip 3328 --launch-count 16 -f -o $OUT/r01_nerf_full \
    python bench.py --steps 4 --warmup 3 > $OUT/ncu_full.log 2>&1
ls -la $OUT/r01_nerf_full.ncu-rep
ncu -i $OUT/r01_nerf_full.ncu-rep --page raw --csv > $OUT/r01_nerf_full_raw.csv 2>/dev/null
echo "raw rows: $(wc -l < $OUT/r01_nerf_full_raw.csv)"
