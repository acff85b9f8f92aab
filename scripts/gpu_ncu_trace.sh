T="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-path"
$T > gpurun_out/ncu_plain_trace.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:trace_kernel -s 3 -c 1 -o gpurun_out/prof_trace -f $T > gpurun_out/ncu_trace2.log 2>&1
tail -3 gpurun_out/ncu_trace2.log
