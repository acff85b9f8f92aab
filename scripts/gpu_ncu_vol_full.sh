# ncu full capture of the volumetric kernels on the fog render (after a plain run of the same command), kernels as plain launches
export MTSB_GRAPH=0
R="python bench_render.py --config vol --no-cpu"
$R > gpurun_out/ncu_plain_vol.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'volshade_kernel|connect_vol_kernel' -s 6 -c 6 -o gpurun_out/prof_vol -f $R > gpurun_out/ncu_vol2.log 2>&1
ls -la gpurun_out/prof_vol.ncu-rep
