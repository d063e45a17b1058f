mkdir -p gpurun_out
python tools/ncu_step.py 256 eval efficientformerv2_s1 > gpurun_out/plain_efv2.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:attn_mma_th -c 1 -o gpurun_out/attn_th -f python tools/ncu_step.py 256 eval efficientformerv2_s1 > gpurun_out/ncu_attn_th.log 2>&1
python tools/ncu_step.py 256 eval faster_vit_0_224 > gpurun_out/plain_fv0.log 2>&1 && \
ncu --set full --clock-control none --profile-from-start off --page raw --csv --log-file gpurun_out/r02_step_raw_fv0.csv python tools/ncu_step.py 256 eval faster_vit_0_224 > gpurun_out/ncu_fv0.log 2>&1
ncu --set full --clock-control none --profile-from-start off --page raw --csv --log-file gpurun_out/r02_step_raw_efv2.csv python tools/ncu_step.py 256 eval efficientformerv2_s1 > gpurun_out/ncu_efv2.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/r02_step_raw*
