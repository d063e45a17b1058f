# round-end evidence on one B200: full GPU suite, smoke, every bench line, the unmodified reference on the backend
export PYTHONPATH=$PWD
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/f_tests.log 2>&1; tail -4 gpurun_out/f_tests.log
python __graft_entry__.py smoke > gpurun_out/f_smoke.log 2>&1; tail -9 gpurun_out/f_smoke.log
python bench.py > gpurun_out/f_default.json 2> gpurun_out/f.err
python bench.py --model efficientformerv2_s1 --no-extras > gpurun_out/f_efv2.json 2>> gpurun_out/f.err
python bench.py --model faster_vit_0_224 --no-extras > gpurun_out/f_fv0.json 2>> gpurun_out/f.err
python bench.py --model faster_vit_2_224 --no-e2e --no-cpu-baseline > gpurun_out/f_fv2.json 2>> gpurun_out/f.err
python bench.py --model efficientnet_b3 --no-e2e --no-cpu-baseline > gpurun_out/f_b3.json 2>> gpurun_out/f.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/f_ref.json 2>> gpurun_out/f.err
tail -3 gpurun_out/f.err
# the reference checkout does not exist on the GPU box: for this one evidence run a temporary, git-ignored copy of it
# (`cp -r /root/reference .refcopy`, deleted again afterwards) travels with the snapshot; nothing in the product reads it
if [ -d .refcopy ]; then
  python tools/run_reference_on_backend.py --reference-root $PWD/.refcopy --ui --images 16 > gpurun_out/f_reference_on_backend.txt 2>&1; tail -14 gpurun_out/f_reference_on_backend.txt
  # the same web_ui run with Grad-CAM through hooks + loss.backward() (pytorch-grad-cam's mechanism) instead of the fused call
  B200DF_GRADCAM_HOOKS=1 python tools/run_reference_on_backend.py --reference-root $PWD/.refcopy --ui --images 8 --models efficientnet_b3,efficientformerv2_s1,faster_vit_2_224,faster_vit_0_224 > gpurun_out/f_reference_on_backend_hooks.txt 2>&1; tail -8 gpurun_out/f_reference_on_backend_hooks.txt
fi
python tools/ncu_step.py 256 x efficientformerv2_s1 > gpurun_out/f_plain_efv2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/f_launches_efv2.csv python tools/ncu_step.py 256 x efficientformerv2_s1 > gpurun_out/f_ncu_efv2.log 2>&1
tail -1 gpurun_out/f_ncu_efv2.log
