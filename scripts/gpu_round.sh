#!/bin/bash
# One GPU-box session: parity tests, micro-benchmarks, encoder variant sweep, bench, ncu.
# Usage (from the repo root on the box): bash scripts/gpu_round.sh <tag> [stages...]
TAG=${1:-r01}; shift
STAGES=${@:-"tests ubench variants bench ncu"}
OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
for st in $STAGES; do case $st in
tests)
  timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log; tail -5 $OUT/pytest_gpu.log;;
ubench)
  timeout 300 ./scripts/ubench_atomics > $OUT/ubench_atomics.json 2>&1; cat $OUT/ubench_atomics.json;;
variants)
  for v in ${VARIANTS:-0 7 8 9 10 11}; do
    DVSOF_ENCODE_VARIANT=$v timeout 600 python scripts/exp_encoder.py cfg2 cfg3 cfg4 >> $OUT/variants.jsonl 2>> $OUT/variants.err
  done; cat $OUT/variants.jsonl;;
bench)
  timeout 900 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; cat $OUT/bench.json; tail -3 $OUT/bench.err
  timeout 300 python bench.py --impl reference --steps 20 --warmup 3 > $OUT/bench_reference.json 2>> $OUT/bench.err; cat $OUT/bench_reference.json;;
ncu)
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_bench_quick.csv \
      python bench.py --quick --steps 3 --warmup 1 > $OUT/ncu_launches.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'encode_|strip' -c 12 -f -o $OUT/ncu_full_cfg2 \
      python bench.py --quick --steps 2 --warmup 1 > $OUT/ncu_full_cfg2.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'encode_v|encode_t' -c 3 -f -o $OUT/ncu_full_cfg3 \
      python scripts/exp_encoder.py cfg3 > $OUT/ncu_full_cfg3.log 2>&1
  ls -la $OUT;;
loss)
  timeout 600 python scripts/exp_loss.py cfg2 cfg3 cfg4 > $OUT/loss.json 2> $OUT/loss.err; cat $OUT/loss.json; tail -3 $OUT/loss.err;;
lossncu)
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'strip|resize|match|outborder|scale_grad' -c 24 -f -o $OUT/ncu_full_loss_cfg4 \
      python scripts/exp_loss.py cfg4 > $OUT/ncu_full_loss_cfg4.log 2>&1; tail -2 $OUT/ncu_full_loss_cfg4.log;;
esac; done
