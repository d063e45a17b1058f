mkdir -p gpurun_out
python tools/ncu_step.py 256 x faster_vit_0_224 > gpurun_out/plain_fv0.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_fv0_vh.csv python tools/ncu_step.py 256 x faster_vit_0_224 > gpurun_out/ncu_fv0_vh.log 2>&1
tail -1 gpurun_out/ncu_fv0_vh.log
