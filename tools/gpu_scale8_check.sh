# one 8-GPU box: the driver's scaling command for N = 8 (weak scaling, one rank per GPU) + the reference arm under the same launcher
export PYTHONPATH=$PWD
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/s8.json 2> gpurun_out/s8.err; tail -c 400 gpurun_out/s8.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29532 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > gpurun_out/s8_ref.json 2>> gpurun_out/s8.err
python - <<'PY'
import json
for f in ('s8', 's8_ref'):
    d = json.loads(open(f'gpurun_out/{f}.json').read().strip().splitlines()[-1])
    print(f, {k: d.get(k) for k in ('value', 'n_gpus', 'ms_per_step', 'compute_ms_per_step', 'gather_ms')}, (d.get('e2e') or {}).get('value'), (d.get('strong') or {}).get('value'))
PY
