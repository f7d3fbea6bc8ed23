# Round-2 profile pass, training step + the small streaming kernels (run through gpurun).
set -x
T="python bench_train.py --batch 128 --steps 1 --warmup 1 --no-cpu-baseline"
$T > gpurun_out/prof_train.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -s 1700 -c 1700 --csv --log-file gpurun_out/r2_launches_train.csv $T > gpurun_out/ncu_t1.log 2>&1
ncu --set full --clock-control none -k regex:"flash_train_bwd_kv|flash_train_bwd_dq|flash_train_fwd" -s 40 -c 3 -o gpurun_out/r2_full_flash_train -f $T > gpurun_out/ncu_t2.log 2>&1
ncu --set full --clock-control none -k regex:"layernorm_vec_kernel|embed_ln_kernel" -s 2 -c 2 -o gpurun_out/r2_full_layernorm -f python profiles/decode_stack_profile.py 256 > gpurun_out/ncu_t3.log 2>&1
ncu --set full --clock-control none -k regex:"relu_bn_bwd_kernel|layernorm_bwd_kernel" -s 20 -c 2 -o gpurun_out/r2_full_train_elementwise -f $T > gpurun_out/ncu_t4.log 2>&1
ls -la gpurun_out | grep -E "r2_full_(flash_train|layernorm|train_el)|r2_launches_train"
