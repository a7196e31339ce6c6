#!/bin/bash
# Same-box GPU baseline at batch sizes that fill the machine: the reference's own kernels
# rebuilt unmodified for sm_100a vs this library's CLI clone vs the reference's harness over this library, on the
# SAME text files, 2 repeats each; raw logs are kept (profiles/r02_refgpu_*.log).
mkdir -p gpurun_out /tmp/refcmp
for case in "susy 500 10 18 500000" "higgs 200 12 28 200000"; do
  set -- $case
  python tools/gen_text_case.py /tmp/refcmp/$1 $2 $3 $4 $5
  for rep in 1 2; do
    echo "== $1 rep $rep: reference kernels (sm_100a rebuild)"; ( time timeout 900 oracle/_ref/Tahoe_ref /tmp/refcmp/$1.model.txt /tmp/refcmp/$1.data.txt ) > gpurun_out/r02_refgpu_$1_reference_$rep.log 2>&1; echo "rc=$?"
    grep -E "Exec.Time|Results are|speedup|not suitable|real" gpurun_out/r02_refgpu_$1_reference_$rep.log
    echo "== $1 rep $rep: tahoe_b200 CLI clone"; ( time timeout 900 tahoe_b200/bin/Tahoe /tmp/refcmp/$1.model.txt /tmp/refcmp/$1.data.txt ) > gpurun_out/r02_refgpu_$1_ours_$rep.log 2>&1; echo "rc=$?"
    grep -E "Exec.Time|Results are|speedup|not suitable|real" gpurun_out/r02_refgpu_$1_ours_$rep.log
  done
  if [ -x oracle/_ref/Tahoe_dropin ]; then
    echo "== $1: reference harness over libtahoe_b200 (drop-in)"; ( time timeout 900 oracle/_ref/Tahoe_dropin /tmp/refcmp/$1.model.txt /tmp/refcmp/$1.data.txt ) > gpurun_out/r02_refgpu_$1_dropin.log 2>&1; echo "rc=$?"
    grep -E "Exec.Time|Results are|speedup|not suitable|FAIL|real" gpurun_out/r02_refgpu_$1_dropin.log
  fi
done
