# ncu evidence for the two halves of the hot path (B200_PROFILING.md recipe). Each ncu run follows a plain run of the same command.
T="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-path"
export MTSB_GRAPH=0   # ncu does not list the kernels inside a conditional graph node: same kernels as plain launches
R1="python bench_render.py --config 1 --no-cpu"
R3="python bench_render.py --config 3 --no-cpu --spp3 8"
$T > gpurun_out/ncu_plain_trace.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_trace.csv $T > gpurun_out/ncu_trace1.log 2>&1
$T > gpurun_out/ncu_plain_trace.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:trace_kernel -s 3 -c 2 -o gpurun_out/prof_trace -f $T > gpurun_out/ncu_trace2.log 2>&1
$R1 > gpurun_out/ncu_plain_render.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_render.csv $R1 > gpurun_out/ncu_render1.log 2>&1
$R3 > gpurun_out/ncu_plain_c3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_c3.csv $R3 > gpurun_out/ncu_c3_1.log 2>&1
$R3 > gpurun_out/ncu_plain_c3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'extend_kernel|shade_kernel|connect_kernel|splat_kernel|generate_kernel' -s 1 -c 8 -o gpurun_out/prof_render -f $R3 > gpurun_out/ncu_render2.log 2>&1
ls -la gpurun_out | tail -5
