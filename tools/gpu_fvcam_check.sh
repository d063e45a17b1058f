mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_fastervit.py tests/test_gpu_efv2.py tests/test_gpu_two_devices.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/t_fvcam.log 2>&1; tail -5 gpurun_out/t_fvcam.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -9 gpurun_out/smoke.log
python bench.py > gpurun_out/b_default.json 2> gpurun_out/b.err; tail -3 gpurun_out/b.err
python - <<'PY'
import json
d = json.loads(open('gpurun_out/b_default.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['cam']['value'])
for k, v in d['models'].items():
    print(k, v['value'], v['cam'])
PY
