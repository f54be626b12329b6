set -u
mkdir -p gpurun_out
S=profiles/probes/sanitize_all.py
timeout 300 python $S > gpurun_out/san_plain.log 2>&1; echo "plain rc=$?" | tee gpurun_out/san_rc.txt
[ "$(tail -1 gpurun_out/san_rc.txt)" = "plain rc=0" ] || { tail -20 gpurun_out/san_plain.log; exit 1; }
for tool in memcheck initcheck synccheck racecheck; do
  t0=$(date +%s)
  timeout 240 compute-sanitizer --tool $tool --error-exitcode 7 --print-limit 20 python $S > gpurun_out/san_$tool.log 2>&1
  echo "$tool rc=$? secs=$(( $(date +%s) - t0 ))" | tee -a gpurun_out/san_rc.txt
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|\[sanitize\] done" gpurun_out/san_$tool.log | tail -3
done
t0=$(date +%s)
timeout 360 compute-sanitizer --tool memcheck --error-exitcode 7 --print-limit 20 python -m pytest tests -m gpu -q -x -p no:cacheprovider --deselect tests/test_gpu_train.py --deselect tests/test_gpu_pipeline.py > gpurun_out/san_memcheck_pytest.log 2>&1
echo "memcheck-pytest rc=$? secs=$(( $(date +%s) - t0 ))" | tee -a gpurun_out/san_rc.txt
tail -5 gpurun_out/san_memcheck_pytest.log
