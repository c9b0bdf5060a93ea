set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2_t7.txt
python tools/profile_stages.py cfg4 3 > gpurun_out/r2_stages7.json 2>&1
python tools/measure_extract_voxel.py cfg2 cfg4 > gpurun_out/r2_extract_voxel7.json 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches7.csv python tools/profile_stages.py cfg4 1 > gpurun_out/r2_ncu_launch7.log 2>&1
