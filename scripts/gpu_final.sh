#!/bin/bash
# Round-end evidence run on the GPU box: the bench without a profiler first, then the ncu launch list of the
# same command and one `--set full` capture per dominant kernel (few launches each: what comes back from the
# box is capped at 64 MiB).  Usage: bash scripts/gpu_final.sh <tag> [bench] [enc] [loss]
TAG=${1:-r02zz}; shift; STAGES=${@:-"bench enc loss"}; OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
if [[ $STAGES == *bench* ]]; then
timeout 600 python bench.py --quick --steps 3 --warmup 3 > $OUT/bench_quick.json 2> $OUT/bench_quick.err; echo "bench --quick rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_bench_quick.csv \
    python bench.py --quick --steps 3 --warmup 3 > $OUT/ncu_launches.log 2>&1; echo "ncu launch list rc=$?"
fi
if [[ $STAGES == *enc* ]]; then
for cfg in cfg3 cfg4; do
  timeout 900 ncu --set full --clock-control none -k regex:'encode_' -c 4 -f -o $OUT/ncu_enc_$cfg \
      python scripts/exp_binned.py $cfg > $OUT/ncu_enc_$cfg.log 2>&1; echo "ncu enc $cfg rc=$?"
done
timeout 900 ncu --set full --clock-control none -k regex:'encode_|voxel_' -c 4 -f -o $OUT/ncu_enc_cfg2 \
    python scripts/exp_encoder.py cfg2 > $OUT/ncu_enc_cfg2.log 2>&1; echo "ncu enc cfg2 rc=$?"
fi
if [[ $STAGES == *loss* ]]; then
for cfg in cfg3 cfg4; do
  DVSOF_EXP_STEPS=3 timeout 900 ncu --set full --clock-control none -k regex:'warp_loss|pyramid|loss_reduce|scale_grad' -c 8 -f \
      -o $OUT/ncu_loss_$cfg python scripts/exp_loss_fused.py $cfg > $OUT/ncu_loss_$cfg.log 2>&1; echo "ncu loss $cfg rc=$?"
done
fi
ls -la $OUT
