#!/bin/bash
# compute-sanitizer over CI-sized GPU tests (race detection hook).
#   tools/sanitize.sh [out_dir] [tools...]      (run on a B200 box: gpurun -- tools/sanitize.sh)
# One log per tool under out_dir; the pytest output of each run next to it.
out=${1:-gpurun_out/san}; shift
tools=${@:-memcheck racecheck synccheck initcheck}
mkdir -p "$out"
SEL="tests/test_gpu_parity.py::test_screenshot_kat tests/test_gpu_parity.py::test_empty_and_tiny_batches \
tests/test_gpu_parity.py::test_golden_vectors_from_the_unmodified_reference \
tests/test_gpu_parity.py::test_random_adversarial_records tests/test_gpu_parity.py::test_streaming_state_carries_over \
tests/test_gpu_parity.py::test_workload_parity[attack_replay_10k-10000] \
tests/test_gpu_parity.py::test_one_pid_across_many_tiles_and_extreme_pids \
tests/test_gpu_exchange.py::test_direct_exchange_matches_oracle_over_two_steps[2-120000-skewed_zipf_chains] \
tests/test_gpu_exchange.py::test_direct_exchange_matches_oracle_over_two_steps[3-120000-skewed_zipf_chains] \
tests/test_gpu_exchange.py::test_empty_slices_and_empty_steps \
tests/test_gpu_checkpoint.py tests/test_zz_no_alloc.py"
for t in $tools; do
  extra=""
  [ "$t" = memcheck ] && extra="--leak-check no"
  [ "$t" = initcheck ] && extra="--track-unused-memory no"
  timeout ${SAN_TIMEOUT:-900} compute-sanitizer --tool $t $extra --print-limit 40 --log-file "$out/$t.log" \
      python -m pytest $SEL -q -x -p no:cacheprovider > "$out/$t.pytest.log" 2>&1
  echo "$t rc=$?" >> "$out/summary.txt"
  grep -h "ERROR SUMMARY\|RACECHECK SUMMARY" "$out/$t.log" | tail -1 >> "$out/summary.txt"
done
cat "$out/summary.txt"
