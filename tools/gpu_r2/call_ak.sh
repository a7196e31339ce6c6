#!/bin/bash
# round 2, call AK: the full GPU suite after the launch-path trimming
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -q -m gpu > gpurun_out/r2ak_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2ak_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
