mkdir -p gpurun_out/$1
python -m pytest tests/test_encoder_gpu.py -m gpu -x -q 2>&1 | tail -2
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:encode_ -c 80 --csv --log-file gpurun_out/$1/binned_launches.csv python scripts/exp_encoder.py cfg3 cfg4 > gpurun_out/$1/log.txt 2>&1; tail -1 gpurun_out/$1/log.txt
python scripts/exp_encoder.py cfg3 cfg4
