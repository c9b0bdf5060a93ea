set -x
python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2_t5.txt
python tools/profile_stages.py cfg4 3 > gpurun_out/r2_stages5.json 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches5.csv python tools/profile_stages.py cfg4 1 > gpurun_out/r2_ncu_launch5.log 2>&1
for k in sdf_row_kernel sdf_col_pack_smem_kernel mc_count_kernel mc_vertex_kernel mc_face_kernel mc_cube_kernel; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$k -c 1 -f -o gpurun_out/r2b_prof_$k python tools/profile_stages.py cfg4 1 > gpurun_out/r2b_ncu_$k.log 2>&1
done
