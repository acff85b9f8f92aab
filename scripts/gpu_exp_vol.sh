# A/B of volshade variants: builds alternative libraries on the box (nvcc is there) and runs the vol bench + parity on each
set -x
run() {
  tag=$1; shift
  touch mitsuba-im_b200/csrc/render.cu
  make -C mitsuba-im_b200/csrc EXTRA="$*" > gpurun_out/expvol_build_$tag.log 2>&1 || { tail -5 gpurun_out/expvol_build_$tag.log; return; }
  grep -A2 "volshade_kernel" mitsuba-im_b200/lib/obj/render.ptxas.log | grep -o "Used [0-9]* registers" | head -1
  python bench_render.py --config vol --no-cpu 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin)['volpath_fog_1024x768_64spp']; print('$tag', round(d['samples_per_s']/1e6,1), 'M samples/s', round(d['ms'],1), 'ms')"
  python -m pytest tests/test_render_gpu.py -m gpu -q -k volpath -s 2>&1 | grep "pixels within\|passed\|failed"
}
run base
run fp32 -DMTSB_VOL_FP32_EXP
run mb4 -DMTSB_VOLSHADE_MIN_BLOCKS=4
run mb5_fp32 -DMTSB_VOLSHADE_MIN_BLOCKS=5 -DMTSB_VOL_FP32_EXP
