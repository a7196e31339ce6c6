#!/bin/bash
# oracle/make_dropin.sh -- the drop-in claim, executable: the REFERENCE's own harness (main.cu +
# BaseTahoeTest.h: its loaders, its CPU predict, its compare_GPU check, its timing loop) running
# THIS library's kernels.  The only edit is the one INTEGRATION.md section 2 describes: the harness
# constructs `b200_forest` (the binding stub over the C ABI, extracted from INTEGRATION.md) instead
# of `dense_adaptive_forest`.  The patched copies live in a scratch directory; nothing of the
# reference enters the repo; the binary goes to oracle/_ref/Tahoe_dropin (git-ignored).
set -e
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"; ROOT="$(dirname "$HERE")"
[ -f "$REF/main.cu" ] || { echo "oracle: $REF absent"; exit 0; }
W=$(mktemp -d /tmp/tahoe_dropin.XXXXXX)
python3 - "$ROOT/INTEGRATION.md" "$W/tahoe_b200_forest.h" <<'PY'
import re, sys
doc = open(sys.argv[1]).read()
open(sys.argv[2], "w").write(re.search(r"```cpp\n(// tahoe_b200_forest\.h.*?)```", doc, re.S).group(1))
PY
# the harness, with the engine type renamed and the stub included after Struct.h
# ... and the feasibility gates (BaseTahoeTest.h:657-679, the reference's 0.8 x 48 KB budget) replaced
# by the library's own answer, as INTEGRATION.md says
GATE='if (tahoe_strategy_applicable(forest->h, loop, ps.num_rows) != TAHOE_OK) { acc[loop] = FLT_MAX; printf("Strategy %d is not suitable for this case.\\n", loop+1); continue; }'
sed -e 's/dense_adaptive_forest/b200_forest/g' \
    -e 's/#include "Struct.h"/#include "Struct.h"\n#include "tahoe_b200_forest.h"/' \
    -e "s/selected_algorithm = loop;/selected_algorithm = loop; $GATE/" \
    -e 's/int max_shm = (int)dev.sharedMemPerBlock \* 0.8;/int max_shm = 1 << 30;/' \
    "$REF/BaseTahoeTest.h" > "$W/BaseTahoeTest.h"
grep -q 'tahoe_b200_forest.h' "$W/BaseTahoeTest.h" || { echo "could not place the stub include"; exit 1; }
grep -q 'tahoe_strategy_applicable' "$W/BaseTahoeTest.h" || { echo "could not place the gate"; exit 1; }
cp "$REF/main.cu" "$W/main.cu"
mkdir -p "$HERE/_ref"
nvcc -include cfloat -include algorithm -I"$W" -I"$REF/inc" -I"$REF" -I"$ROOT/include" -O3 --expt-extended-lambda \
     -std=c++17 -w -gencode arch=compute_100a,code=sm_100a "$W/main.cu" -o "$HERE/_ref/Tahoe_dropin" \
     -L"$ROOT/tahoe_b200/lib" -ltahoe_b200 -Xlinker -rpath -Xlinker '$ORIGIN/../../tahoe_b200/lib'
rm -rf "$W"
echo "$HERE/_ref/Tahoe_dropin"
