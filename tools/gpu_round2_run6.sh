set -x
python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_shim_demo.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2_t6.txt
python tools/profile_stages.py cfg4 3 > gpurun_out/r2_stages6.json 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches6.csv python tools/profile_stages.py cfg4 1 > gpurun_out/r2_ncu_launch6.log 2>&1
