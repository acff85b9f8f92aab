# launch list of the volpath_simple render (after a plain run of the same command), kernels as plain launches
export MTSB_GRAPH=0
R="python bench_render.py --config vol --no-cpu"
$R > gpurun_out/ncu_plain_vol.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_vol.csv $R > gpurun_out/ncu_vol1.log 2>&1
python scripts/exp/launch_table.py gpurun_out/launches_vol.csv > gpurun_out/launches_vol_table.txt 2>&1
cat gpurun_out/launches_vol_table.txt
