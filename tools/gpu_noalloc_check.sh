mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/t_noalloc_alone.log 2>&1; tail -3 gpurun_out/t_noalloc_alone.log
python -m pytest tests -x -q -m gpu > gpurun_out/t_all.log 2>&1; tail -5 gpurun_out/t_all.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
