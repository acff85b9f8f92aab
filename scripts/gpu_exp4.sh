mkdir -p gpurun_out
# responsive CLI path on config 1 with the per-thread timing log
MTSB_PLUGIN_TIMING=1 oracle/_ref/mitsuba -p 16 -a mitsuba-im_b200/plugin/overlay -D sampler=independent -D spp=64 -D res=512 -D rfilter=box -o /tmp/r.pfm tests/golden/cbox_config1.xml 2>&1 | grep -E "gpupath|Render|SPP" | cut -c1-260
# C3 launch list (8 spp)
R3="python bench_render.py --config 3 --no-cpu --spp3 8"
$R3 > gpurun_out/ncu_plain_c3.log 2>&1 &&
MTSB_GRAPH=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02g_launches_c3.csv $R3 > gpurun_out/ncu_c3_1.log 2>&1
python scripts/exp/launch_table.py gpurun_out/r02g_launches_c3.csv | grep -v "ploc\|cub\|collapse\|morton\|bounds\|triaccel\|record\|init_scene"
