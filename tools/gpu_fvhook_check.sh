mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_fastervit.py tests/test_gpu_efv2.py tests/test_gpu_two_devices.py -x -q -m gpu > gpurun_out/t_hook.log 2>&1; tail -25 gpurun_out/t_hook.log
