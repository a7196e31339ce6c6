#!/bin/bash
# Same-box GPU baseline: the reference's own kernels rebuilt unmodified for sm_100a
# (oracle/_ref/Tahoe_ref, `make -C oracle ref_gpu`) and this library's CLI clone on the SAME text files;
# third arm: the reference's OWN harness running this library's kernels through the INTEGRATION.md
# binding stub (oracle/_ref/Tahoe_dropin, `bash oracle/make_dropin.sh`).
mkdir -p gpurun_out /tmp/refcmp
for case in "susy 500 10 18 50000" "svhn 100 8 3072 3000" "higgs 200 12 28 20000"; do
  set -- $case
  python tools/gen_text_case.py /tmp/refcmp/$1 $2 $3 $4 $5
  echo "== $1: reference kernels (sm_100a rebuild)"; ( time timeout 600 oracle/_ref/Tahoe_ref /tmp/refcmp/$1.model.txt /tmp/refcmp/$1.data.txt ) > gpurun_out/refgpu_$1_reference.log 2>&1; echo "rc=$?"
  grep -E "Exec.Time|Results are|speedup|not suitable|real" gpurun_out/refgpu_$1_reference.log
  echo "== $1: tahoe_b200 CLI clone"; ( time timeout 600 tahoe_b200/bin/Tahoe /tmp/refcmp/$1.model.txt /tmp/refcmp/$1.data.txt ) > gpurun_out/refgpu_$1_ours.log 2>&1; echo "rc=$?"
  grep -E "Exec.Time|Results are|speedup|not suitable|real" gpurun_out/refgpu_$1_ours.log
  if [ -x oracle/_ref/Tahoe_dropin ]; then
    echo "== $1: reference harness over libtahoe_b200 (drop-in)"; ( time timeout 600 oracle/_ref/Tahoe_dropin /tmp/refcmp/$1.model.txt /tmp/refcmp/$1.data.txt ) > gpurun_out/refgpu_$1_dropin.log 2>&1; echo "rc=$?"
    grep -E "Exec.Time|Results are|speedup|not suitable|FAIL|real" gpurun_out/refgpu_$1_dropin.log
  fi
done
