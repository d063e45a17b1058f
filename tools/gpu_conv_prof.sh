# ncu --set full of the conv3x3 launches of one FasterViT-0 forward (read with: ncu -i gpurun_out/prof_conv.ncu-rep --page raw --csv)
mkdir -p gpurun_out
python tools/ncu_step.py 256 x faster_vit_0_224 > gpurun_out/plain_fv0.log 2>&1 && ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:conv3x3_tc -s 1 -c 8 -o gpurun_out/prof_conv python tools/ncu_step.py 256 x faster_vit_0_224 > gpurun_out/ncu_conv.log 2>&1
tail -2 gpurun_out/ncu_conv.log
