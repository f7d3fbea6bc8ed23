python scratch/enc_only.py || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"gemm_tcgen05|stem_im2col|maxpool" --csv --log-file gpurun_out/r1_backbone_v2_raw.csv python scratch/enc_only.py > gpurun_out/ncu_bb.log 2>&1
tail -2 gpurun_out/ncu_bb.log
