cp mitsuba-im_b200/lib/libmtsb200.so /tmp/orig.so
for v in orig sm6 sm8; do
  if [ $v = orig ]; then cp /tmp/orig.so mitsuba-im_b200/lib/libmtsb200.so; else cp variants/$v/libmtsb200.so mitsuba-im_b200/lib/libmtsb200.so; fi
  echo "== $v"
  timeout 600 python bench_render.py --config 1 3 --no-cpu --spp3 64 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read())
for k,v in d.items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ('samples_per_s','ms')})"
done
cp /tmp/orig.so mitsuba-im_b200/lib/libmtsb200.so
