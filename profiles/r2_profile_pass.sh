# Round-2 profile pass (run on the GPU box through gpurun; reports land in gpurun_out/, summaries are copied to profiles/).
# Every ncu command re-runs a program that has already exited 0 without ncu.
set -x
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-train-leg --no-parity-leg"
$B > gpurun_out/prof_bench.log 2>&1 || exit 1
# (1) launch list of one whole step: backbone + encoder (first ~330 launches), then decode steps (3 launches each)
ncu --metrics gpu__time_duration.sum --clock-control none -c 420 --csv --log-file gpurun_out/r2_launches_front.csv $B > gpurun_out/ncu1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"decode_stack|greedy_update|gemm_tcgen05" -s 600 -c 90 --csv --log-file gpurun_out/r2_launches_decode.csv $B > gpurun_out/ncu2.log 2>&1
# (2) full captures: decode stack kernel (mid-sequence step), encoder flash attention, vocab GEMM + argmax, LayerNorm
ncu --set full --clock-control none --import-source on -k regex:decode_stack_kernel -s 60 -c 1 -o gpurun_out/r2_full_decode_stack -f $B > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:flash_attention_kernel -s 2 -c 1 -o gpurun_out/r2_full_flash_attention -f $B > gpurun_out/ncu4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:layernorm_kernel -s 2 -c 1 -o gpurun_out/r2_full_layernorm -f $B > gpurun_out/ncu5.log 2>&1
# the vocabulary GEMM of a decode step is the gemm_tcgen05 launch right after a decode_stack launch: grid 148, N = 30522
ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -s 330 -c 2 -o gpurun_out/r2_full_vocab_gemm -f $B > gpurun_out/ncu6.log 2>&1
# (3) DRAM traffic of the backbone phase (bench.py roofline.traffic)
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"gemm_tcgen05|stem_im2col|maxpool" -c 110 --csv --log-file gpurun_out/r2_backbone_dram.csv $B > gpurun_out/ncu7.log 2>&1
# (4) preprocessing kernel
python profiles/preprocess_bench.py > gpurun_out/r2_preprocess_bench.json 2> gpurun_out/pp.err
ncu --set full --clock-control none --import-source on -k regex:preprocess_kernel -s 2 -c 1 -o gpurun_out/r2_full_preprocess -f python profiles/preprocess_bench.py > gpurun_out/ncu8.log 2>&1
ls -la gpurun_out | tail -20
