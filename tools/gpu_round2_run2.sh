set -x
python -m pytest tests/test_gpu_sharded.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2_t2.txt
python bench.py --steps 3 --warmup 3 > gpurun_out/r2_b2.json 2> gpurun_out/r2_b2.err
python bench.py --steps 2 --warmup 3 --workload cfg4_nomask --no-cpu-baseline > gpurun_out/r2_nomask.json 2> gpurun_out/r2_nomask.err
python bench.py --steps 3 --warmup 3 --workload cfg3 --no-cpu-baseline > gpurun_out/r2_cfg3.json 2> gpurun_out/r2_cfg3.err
python bench.py --steps 3 --warmup 3 --workload cfg2 --no-cpu-baseline > gpurun_out/r2_cfg2.json 2> gpurun_out/r2_cfg2.err
python tools/profile_stages.py cfg4 3 > gpurun_out/r2_stages_cfg4.json 2>&1
python tools/measure_extract_voxel.py cfg2 cfg4 > gpurun_out/r2_extract_voxel.json 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches_cfg4.csv python tools/profile_stages.py cfg4 1 > gpurun_out/r2_ncu_launch.log 2>&1
for k in sdf_bits_kernel sdf_row_kernel sdf_col_pack_kernel sdf_lut_kernel sdf_finalize_kernel mc_bits_kernel mc_cube_kernel mc_count_kernel mc_vertex_kernel mc_face_kernel voxel_flag_kernel; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:^$k\$ -c 1 -f -o gpurun_out/r2_prof_$k python tools/profile_stages.py cfg4 1 > gpurun_out/r2_ncu_$k.log 2>&1
done
ls -la gpurun_out | tail -20
