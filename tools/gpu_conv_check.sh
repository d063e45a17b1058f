# A/B of the dense 3x3 conv schedules + the kernels that changed with them (profiles/r02_d_conv3x3.txt): B200DF_CONV_VHALO=0 is the one-box-per-tap schedule
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_vit_kernels.py -x -q -m gpu -k "conv3x3 or layernorm or upsample" > gpurun_out/t_conv.log 2>&1; tail -6 gpurun_out/t_conv.log
timeout 900 python -m pytest tests/test_gpu_fastervit.py tests/test_gpu_efv2.py -x -q -m gpu > gpurun_out/t_conv2.log 2>&1; tail -4 gpurun_out/t_conv2.log
for m in faster_vit_0_224 faster_vit_2_224; do
python bench.py --model $m --no-extras --no-cpu-baseline --no-e2e > gpurun_out/b_vh1_$m.json 2> gpurun_out/b_vh.err
B200DF_CONV_VHALO=0 python bench.py --model $m --no-extras --no-cpu-baseline --no-e2e > gpurun_out/b_vh0_$m.json 2>> gpurun_out/b_vh.err
done
python bench.py --model efficientformerv2_s1 --no-extras --no-cpu-baseline --no-e2e > gpurun_out/b_vh_efv2.json 2>> gpurun_out/b_vh.err
tail -2 gpurun_out/b_vh.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/b_vh*.json')):
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f, d['value'], d['ms_per_step'], {k: (v['ms'], v['launches'], v.get('TFLOPs')) for k, v in d['kernels'].items() if k in ('conv3x3', 'layernorm', 'upsample', 'attention')})
PY
