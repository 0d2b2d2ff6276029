# compute-sanitizer over the small GPU cases (memcheck: out-of-bounds / misaligned accesses; racecheck: shared-memory hazards of the
# permutation table, the probe kernels' sorts and reductions).  gpurun --timeout 1500 -- 'bash tools/sanitize.sh'
S=/usr/local/cuda/bin/compute-sanitizer
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > /dev/null 2>&1 || { echo "smoke failed without the sanitizer"; exit 1; }
echo "== memcheck: smoke()"
timeout 600 $S --tool memcheck --error-exitcode 9 --print-limit 5 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep -E "ERROR SUMMARY|Invalid|out of bounds|misaligned|=========     at" | head -12; echo "rc=${PIPESTATUS[0]}"
echo "== memcheck: small parity / edge / probe / no-alloc tests"
timeout 1200 $S --tool memcheck --error-exitcode 9 --print-limit 5 python -m pytest tests/test_gpu_parity.py tests/test_probes.py tests/test_zz_no_alloc.py -m gpu -x -q -k "not full and not atrium" 2>&1 | grep -E "ERROR SUMMARY|Invalid|passed|failed|error" | head -12; echo "rc=${PIPESTATUS[0]}"
echo "== racecheck: smoke() and the probe tests"
timeout 600 $S --tool racecheck --error-exitcode 9 --print-limit 5 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep -E "RACECHECK SUMMARY|hazard|ERROR" | head -8; echo "rc=${PIPESTATUS[0]}"
timeout 900 $S --tool racecheck --error-exitcode 9 --print-limit 5 python -m pytest tests/test_probes.py -m gpu -x -q 2>&1 | grep -E "RACECHECK SUMMARY|hazard|passed|failed|ERROR" | head -8; echo "rc=${PIPESTATUS[0]}"
