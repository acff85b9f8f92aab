# ncu full capture of the mesh-file decode kernels on the config-3 meshes (after a plain run of the same command)
R="python bench_render.py --config ply --no-cpu"
$R > gpurun_out/ncu_plain_ply.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'ply_vertex_kernel|ply_face_kernel|ply_normalize_kernel' -c 6 -o gpurun_out/prof_ply -f $R > gpurun_out/ncu_ply.log 2>&1
ls -la gpurun_out | tail -4
