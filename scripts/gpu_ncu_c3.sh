R="python bench_render.py --config 3 --no-cpu --spp3 8"
$R > gpurun_out/ncu_plain_c3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_c3.csv $R > gpurun_out/ncu_c3_1.log 2>&1
$R > gpurun_out/ncu_plain_c3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'extend_kernel|shade_kernel|connect_kernel' -s 130 -c 9 -o gpurun_out/prof_c3 -f $R > gpurun_out/ncu_c3_2.log 2>&1
tail -2 gpurun_out/ncu_c3_2.log
