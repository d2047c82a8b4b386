# compute-sanitizer over one small problem per kernel family (run under gpurun; output -> gpurun_out/sanitize_*.txt, the summaries
# are copied to profiles/).  memcheck on everything; racecheck (shared-memory hazards) and synccheck per family with a time limit:
# the tools serialise the kernels heavily.
set -u
python tests/sanitize_driver.py > gpurun_out/sanitize_plain.txt 2>&1 || { echo "driver fails without the sanitizer"; tail -5 gpurun_out/sanitize_plain.txt; exit 1; }
for tool in memcheck racecheck synccheck; do
  for fam in gemv2 gemv fused w8 i8; do
    timeout 420 compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 python tests/sanitize_driver.py $fam > gpurun_out/sanitize_${tool}_${fam}.txt 2>&1
    rc=$?
    echo "$tool $fam rc=$rc $(grep -c 'ERROR SUMMARY' gpurun_out/sanitize_${tool}_${fam}.txt) $(grep 'ERROR SUMMARY\|RACECHECK SUMMARY\|sanitize driver' gpurun_out/sanitize_${tool}_${fam}.txt | tr '\n' ' ')"
  done
done
