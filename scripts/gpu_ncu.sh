# ncu evidence for the two halves of the hot path (B200_PROFILING.md recipe). Each ncu run follows a plain run of the same command.
set -x
T="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-path"
R="python bench_render.py --config 1 --no-cpu"
$T > gpurun_out/ncu_plain_trace.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_trace.csv $T > gpurun_out/ncu_trace1.log 2>&1
$T > gpurun_out/ncu_plain_trace.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:trace_kernel -s 3 -c 2 -o gpurun_out/prof_trace -f $T > gpurun_out/ncu_trace2.log 2>&1
$R > gpurun_out/ncu_plain_render.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_render.csv $R > gpurun_out/ncu_render1.log 2>&1
$R > gpurun_out/ncu_plain_render.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'extend_kernel|shade_kernel|connect_kernel' -s 24 -c 6 -o gpurun_out/prof_render -f $R > gpurun_out/ncu_render2.log 2>&1
ls -la gpurun_out
