set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/check_pytest.txt
python bench_train.py 2>&1 | tail -1 > gpurun_out/check_train.json; cat gpurun_out/check_train.json
# full capture of the TMA-im2col convolution kernel (layer3 conv2 shape is launch ~60 of an encode) and the c3 1x1 + residual kernel
ncu --set full --clock-control none --import-source on -k regex:"gemm_tcgen05_kernel<256, 2>" -s 30 -c 1 -o gpurun_out/r1_full_conv_im2col -f python scratch/enc_only.py > gpurun_out/ncu_full1.log 2>&1
tail -2 gpurun_out/ncu_full1.log
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"gemm_tcgen05|stem_im2col|maxpool" --csv --log-file gpurun_out/r1_backbone_v2_dram.csv python scratch/enc_only.py > gpurun_out/ncu_dram.log 2>&1
ls -la gpurun_out | tail -8
