#!/bin/bash
# round 2, call M: shared-memory records for the internal nodes at the sparse tops' bottom level (config 4)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py tests/test_zz_no_alloc.py tests/test_importers.py -m gpu -x -q > gpurun_out/r2m_pytest_sparse.log 2>&1; rc=$?; echo "pytest(sparse paths) rc=$rc"; tail -12 gpurun_out/r2m_pytest_sparse.log | cut -c1-200
L=gpurun_out/r2m_qtop.log; : > $L
timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1
for d in 4 5 6 7; do TAHOE_QTOP_DEPTH=$d timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1; done
cut -c1-260 $L
if [ $rc -eq 0 ]; then
  TAHOE_FUZZ_SOAK=8 timeout 1500 python -m pytest tests/test_gpu_fuzz.py -x -q > gpurun_out/r2m_fuzz_soak.log 2>&1; echo "fuzz soak rc=$?"; tail -3 gpurun_out/r2m_fuzz_soak.log
  timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2m_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2m_pytest.log
fi
