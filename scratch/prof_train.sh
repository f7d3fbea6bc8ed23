ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r1_train_v2_raw.csv python bench_train.py --batch 128 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_train.log 2>&1
tail -2 gpurun_out/ncu_train.log
# full capture of one TMA-im2col convolution (layer3 conv2 shape: 30th tcgen05 launch of an encode)
ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05_kernel -s 29 -c 1 -o gpurun_out/r1_full_conv_im2col -f python scratch/enc_only.py > gpurun_out/ncu_full1.log 2>&1
tail -2 gpurun_out/ncu_full1.log
