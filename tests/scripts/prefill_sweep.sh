# Device time of the prefill-shaped weight-quantised GEMMs (eager launches, rotating operands > L2)
S="w8 4096 14336 4096 w4 4096 11008 4096 w4 4096 4096 11008 w4 4096 28672 8192 w4 4096 4096 4096 w8 2048 4096 4096"
echo "== fused"; python tests/perf_probe.py $S
if [ -n "$TWO_PASS" ]; then echo "== two-pass (round 1)"; CKB_PREFILL_TWO_PASS=1 python tests/perf_probe.py $S; fi
