# Round-2 closing profile pass (gpurun; reports land in gpurun_out/, summaries are copied to profiles/ by
# profiles/r2_profile_post_final.sh).  Every ncu command re-runs a program that has already exited 0 without ncu.
set -x
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-train-leg --no-parity-leg"
$B > gpurun_out/prof_bench.log 2>&1 || exit 1
# (1) launch list of the front of a step: backbone + encoder + the first decode steps (2 launches each)
ncu --metrics gpu__time_duration.sum --clock-control none -c 420 --csv --log-file gpurun_out/r2f_launches_front.csv $B > gpurun_out/ncu1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"decode_stack|vocab_argmax" -s 100 -c 60 --csv --log-file gpurun_out/r2f_launches_decode.csv $B > gpurun_out/ncu2.log 2>&1
# (2) full captures: the vocabulary slice kernel (mid-sequence step), layer3 block 1 convolutions (conv1, conv2, conv3)
ncu --set full --clock-control none --import-source on -k regex:vocab_argmax_kernel -s 60 -c 1 -o gpurun_out/r2f_full_vocab_argmax -f $B > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -s 28 -c 3 -o gpurun_out/r2f_full_layer3_convs -f $B > gpurun_out/ncu4.log 2>&1
# (3) DRAM traffic of the backbone phase (bench.py roofline_backbone.traffic)
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"gemm_tcgen05|stem_im2col|stem_fused|maxpool" -c 110 --csv --log-file gpurun_out/r2f_backbone_dram.csv $B > gpurun_out/ncu5.log 2>&1
# (4) decode stack kernel at a mid-sequence step (bench.py roofline.traffic) and the persistent fused stem
ncu --set full --clock-control none --import-source on -k regex:decode_stack_kernel -s 60 -c 1 -o gpurun_out/r2f_full_decode_stack -f $B > gpurun_out/ncu6.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:stem_fused_kernel -s 1 -c 1 -o gpurun_out/r2f_full_stem_fused_v3 -f $B > gpurun_out/ncu7.log 2>&1
for r in r2f_full_vocab_argmax r2f_full_layer3_convs r2f_full_decode_stack r2f_full_stem_fused_v3; do
  ncu -i gpurun_out/$r.ncu-rep --page raw --csv > gpurun_out/$r.raw.csv 2>/dev/null
done
ls -la gpurun_out | tail -12
