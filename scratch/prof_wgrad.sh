python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"gemm_tcgen05_kernel<256, 4>" -s 40 -c 1 -o gpurun_out/r1_full_wgrad_im2col -f python bench_train.py --batch 128 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_wgrad.log 2>&1
tail -3 gpurun_out/ncu_wgrad.log
