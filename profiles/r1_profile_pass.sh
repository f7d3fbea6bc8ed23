set -x
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/prof_bench.log 2>&1 || exit 1
tail -c 600 gpurun_out/prof_bench.log
# launch list: backbone + encoder of the first (warm-up) step, then a window of decode steps
ncu --metrics gpu__time_duration.sum --clock-control none -c 420 --csv --log-file gpurun_out/r1_launches_front.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -s 2000 -c 1400 --csv --log-file gpurun_out/r1_launches_decode.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
for k in ffn_fused_kernel attn_out_ln_proj_kernel decode_attention_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 40 -c 2 -o gpurun_out/r1_full_$k -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_$k.log 2>&1
done
ls -la gpurun_out
# DRAM traffic of the backbone phase (roofline.traffic in bench.py): every conv / stem / pool launch of one step
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"gemm_tcgen05|stem_im2col|maxpool" -c 120 --csv --log-file gpurun_out/r1_backbone_dram.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu3.log 2>&1
