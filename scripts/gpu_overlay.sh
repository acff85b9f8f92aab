# per-ray overlay: tests + throughput of the stock CPU `path` integrator with and without it
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_overlay_gpu.py tests/test_plugin_gpu.py -m gpu -q -rA -s 2>&1 | grep -E "overlay \(|passed|failed|^FAILED|^E  " | cut -c1-400 | head -30
M=oracle/_ref/mitsuba; X=tests/golden/cbox_config1.xml; D="-D sampler=independent -D spp=16 -D res=256 -D rfilter=box"
for p in 16 64 256 1024; do
  a=$($M -C -p $p $D -o /tmp/a.pfm $X 2>&1 | grep -E "Render time|Normal rays|Shadow rays" | sed 's/.*\] //; s/^ *- *//' | tr '\n' ';')
  b=$(LD_PRELOAD=mitsuba-im_b200/plugin/overlay/lib/libmtsb_overlay.so $M -C -p $p $D -o /tmp/b.pfm $X 2>&1 | grep -E "Render time|queue:|Closest-hit rays|Shadow rays" | sed 's/.*\] //; s/^ *- *//' | tr '\n' ';')
  echo "threads=$p | stock: $a | overlay: $b"
done 2>&1 | tee gpurun_out/r02_overlay_throughput.log
