# one FasterViT-0 forward (N = 256, bf16) under ncu --set full -> gpurun_out/step_raw_fv0.csv (summarise: tools/summarize_ncu.py step)
mkdir -p gpurun_out
python tools/ncu_step.py 256 x faster_vit_0_224 > gpurun_out/plain_fv0.log 2>&1 && ncu --set full --clock-control none --profile-from-start off --page raw --csv --log-file gpurun_out/step_raw_fv0.csv python tools/ncu_step.py 256 x faster_vit_0_224 > gpurun_out/ncu_step_fv0.log 2>&1
tail -1 gpurun_out/ncu_step_fv0.log
