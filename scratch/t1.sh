python -m pytest tests/test_model_gpu.py -x -q -k "stream or eval_model or generate_host" 2>&1 | tail -3
python bench_train.py --batch 128 --steps 1 --warmup 1 --no-cpu-baseline 2>&1 | tail -1 | cut -c1-400 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -s 2000 -c 4200 --csv --log-file gpurun_out/r1_train_v2_raw.csv python bench_train.py --batch 128 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_train.log 2>&1
tail -2 gpurun_out/ncu_train.log | cut -c1-300
