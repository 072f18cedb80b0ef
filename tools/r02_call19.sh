#!/bin/bash
# round 2, GPU call 19: keep_records (records that survive a launch boundary), forward_many: full suite, smoke, adapter-level timing of
# feature_pnp_multi rounds; compute-sanitizer memcheck on the small all-paths case
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_c19_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_c19_pytest.log
tail -n 3 gpurun_out/r02_c19_pytest.log
python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/r02_c19_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r02_c19_smoke.log
tail -n 2 gpurun_out/r02_c19_smoke.log
python tools/keep_records_rounds.py > gpurun_out/r02_c19_keep_records.txt 2>&1
cat gpurun_out/r02_c19_keep_records.txt
timeout 900 compute-sanitizer --tool memcheck python tools/sanitize_case.py > gpurun_out/r02_c19_sanitize_memcheck.log 2>&1
echo "sanitizer rc=$?" >> gpurun_out/r02_c19_sanitize_memcheck.log
tail -n 8 gpurun_out/r02_c19_sanitize_memcheck.log
