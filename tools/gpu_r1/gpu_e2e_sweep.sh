for cfg in "2,150,16" "2,200,8" "2,150,12" "3,150,16" "2,130,16" "2,150,24" "4,100,4" "8,100,8"; do
  echo "== $cfg"; TAHOE_HOST_SLAB=$cfg timeout 200 python tools/probe_e2e.py 2>&1 | grep predict_host
done
