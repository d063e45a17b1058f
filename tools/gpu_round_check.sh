mkdir -p gpurun_out
python -m pytest tests/test_gpu_vit_kernels.py tests/test_gpu_efv2.py tests/test_gpu_fastervit.py -q -m gpu -k "attention or efv2 or fastervit" > gpurun_out/t4.log 2>&1; tail -4 gpurun_out/t4.log
python __graft_entry__.py smoke > gpurun_out/smoke4.log 2>&1; tail -9 gpurun_out/smoke4.log
python bench.py > gpurun_out/b4_default.json 2> gpurun_out/b4.err
python bench.py --model efficientformerv2_s1 --no-extras > gpurun_out/b4_efv2.json 2>> gpurun_out/b4.err
python bench.py --model faster_vit_0_224 --no-extras > gpurun_out/b4_fv0.json 2>> gpurun_out/b4.err
python bench.py --model faster_vit_2_224 --no-e2e --no-cpu-baseline > gpurun_out/b4_fv2.json 2>> gpurun_out/b4.err
python bench.py --model efficientnet_b3 --no-e2e --no-cpu-baseline > gpurun_out/b4_b3.json 2>> gpurun_out/b4.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/b4_ref.json 2>> gpurun_out/b4.err
tail -3 gpurun_out/b4.err
python tools/ncu_step.py 256 > gpurun_out/ncu_plain.log 2>&1 && ncu --set full --clock-control none --profile-from-start off --page raw --csv --log-file gpurun_out/r02_step_raw.csv python tools/ncu_step.py 256 > gpurun_out/ncu4.log 2>&1
tail -2 gpurun_out/ncu4.log
