mkdir -p gpurun_out
python tools/ncu_step.py 256 eval faster_vit_0_224 > gpurun_out/plain_fv0.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:attn_mma -c 3 -o gpurun_out/attn_fv0 -f python tools/ncu_step.py 256 eval faster_vit_0_224 > gpurun_out/ncu_attn_fv0.log 2>&1
python tools/ncu_step.py 256 eval efficientformerv2_s1 > gpurun_out/plain_efv2.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:attn_mma -c 3 -o gpurun_out/attn_efv2 -f python tools/ncu_step.py 256 eval efficientformerv2_s1 > gpurun_out/ncu_attn_efv2.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02_launches_fv0.csv python tools/ncu_step.py 256 eval faster_vit_0_224 > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02_launches_efv2.csv python tools/ncu_step.py 256 eval efficientformerv2_s1 > /dev/null 2>&1
ls -la gpurun_out/*.ncu-rep
