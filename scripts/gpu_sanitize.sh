# compute-sanitizer (memcheck, racecheck, synccheck) over the smoke-sized trace + render; logs -> gpurun_out/ (copied to profiles/)
mkdir -p gpurun_out
S="python -c 'import __graft_entry__ as g; g.smoke()'"
for tool in memcheck racecheck synccheck initcheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 bash -c "$S" > gpurun_out/sanitizer_$tool.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|smoke" gpurun_out/sanitizer_$tool.log | tail -4
done
