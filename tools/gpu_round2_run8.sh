set -x
python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2_t8b.txt
python tools/profile_stages.py cfg4 3 > gpurun_out/r2_stages8.json 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches8.csv python tools/profile_stages.py cfg4 1 > gpurun_out/r2_ncu_launch8.log 2>&1
for k in mc_cube_kernel mc_count_kernel mc_vertex_kernel mc_face_kernel mc_rowprefix_kernel voxel_flag_kernel sdf_finalize_kernel; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$k -c 1 -f -o gpurun_out/r2d_prof_$k python tools/profile_stages.py cfg4 1 > gpurun_out/r2d_ncu_$k.log 2>&1
done
