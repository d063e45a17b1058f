# 2-GPU box: two-device test, the unmodified reference on the backend, N = 2 bench + reference arm, sharded harness
export PYTHONPATH=$PWD
mkdir -p gpurun_out
python -m pytest tests/test_gpu_two_devices.py -q -m gpu > gpurun_out/t9_two_dev.log 2>&1; tail -3 gpurun_out/t9_two_dev.log
if [ -d .refcopy ]; then
  python tools/run_reference_on_backend.py --reference-root $PWD/.refcopy --ui --images 16 > gpurun_out/r02_reference_on_backend.txt 2>&1; tail -12 gpurun_out/r02_reference_on_backend.txt
fi
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/b9_n2.json 2> gpurun_out/b9_n2.err; tail -c 300 gpurun_out/b9_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/b9_ref_n2.json 2>> gpurun_out/b9_n2.err
W=/tmp/harness_demo; python tools/harness_check.py prepare $W > gpurun_out/t9_harness.log 2>&1
(cd $W && python -m deepfakedetection_b200.harness --config inference.yaml >> $OLDPWD/gpurun_out/t9_harness.log 2>&1; mkdir -p one && cat runs/*/*/logs/metrics.jsonl > one/metrics.jsonl; rm -rf runs)
(cd $W && PYTHONPATH=$OLDPWD python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 -m deepfakedetection_b200.harness --config inference.yaml >> $OLDPWD/gpurun_out/t9_harness.log 2>&1; mkdir -p two && cat runs/*/*/logs/metrics.jsonl > two/metrics.jsonl)
python tools/harness_check.py compare $W/one/metrics.jsonl $W/two/metrics.jsonl >> gpurun_out/t9_harness.log 2>&1; tail -4 gpurun_out/t9_harness.log
