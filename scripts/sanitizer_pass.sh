# compute-sanitizer over the small-shape decode tests that exercise the shared-memory segment queue,
# the deferred acq_rel merge arrivals and the self re-arming counters.
# Logs go to gpurun_out/; the tails are committed under profiles/.
SEL='minimum_slice or split_kv or many_short or empty_request or fp8_kv'
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --target-processes all --print-limit 20 \
    python -m pytest tests/test_decode_gpu.py -m gpu -x -q -k "$SEL" > gpurun_out/sanitizer_$tool.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" gpurun_out/sanitizer_$tool.log | tail -5
done
