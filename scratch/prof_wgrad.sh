# full capture of the in-place convolution weight gradient (MN-major operands, TMA im2col B): one launch of layer3
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"gemm_tcgen05_kernel<.int.256, .int.4>" -s 40 -c 1 -o gpurun_out/r1_full_wgrad_im2col -f python bench_train.py --batch 128 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_wgrad.log 2>&1
tail -2 gpurun_out/ncu_wgrad.log
