#!/bin/bash
# round 2, call 8t: sanity of the final library (after dead-code removal and comment edits): smoke + the Add parity files
mkdir -p gpurun_out
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r8t_smoke.log 2>&1; tail -2 gpurun_out/r8t_smoke.log | cut -c1-200
timeout 80 python -m pytest tests/test_add_window_gpu.py tests/test_search_parity_gpu.py -q -m gpu -x > gpurun_out/r8t_pytest.log 2>&1
echo "tests rc=$?"; tail -2 gpurun_out/r8t_pytest.log | cut -c1-200
