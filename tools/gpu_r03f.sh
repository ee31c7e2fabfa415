#!/bin/bash
# r03f (8 GPUs): why does ark -> x-vector ark get SLOWER with more GPUs?  ring size and batch size sweeps of the ark_e2e leg
out=gpurun_out; tag=r03f
cat > /tmp/ark_leg.py <<'PY'
import json, os, sys
sys.path.insert(0, os.getcwd())
import bench
class A: pass
n = int(sys.argv[1])
r = bench.leg_ark_e2e(A(), list(range(n)))
print(json.dumps({"devices": n, "ring": os.environ.get("XV_EXTRACT_RING"), "rows": os.environ.get("XV_ARK_ROWS"),
                  "fm": round(r["fm"]["frames_per_s"] / 1e6, 1), "cm": round(r["cm"]["frames_per_s"] / 1e6, 1),
                  "reader_fm": round(r["fm"]["reader_alone_frames_per_s"] / 1e6, 1)}))
PY
for n in 1 2 8; do python /tmp/ark_leg.py $n 2>/dev/null | tail -1; done
for ring in 3 5 9; do XV_EXTRACT_RING=$ring python /tmp/ark_leg.py 8 2>/dev/null | tail -1; done
for rows in 131072 262144; do XV_ARK_ROWS=$rows python /tmp/ark_leg.py 8 2>/dev/null | tail -1; done
XV_EXTRACT_RING=5 XV_ARK_ROWS=131072 python /tmp/ark_leg.py 8 2>/dev/null | tail -1
XV_EXTRACT_RING=5 XV_ARK_ROWS=131072 python /tmp/ark_leg.py 2 2>/dev/null | tail -1
XV_EXTRACT_RING=5 XV_ARK_ROWS=131072 python /tmp/ark_leg.py 1 2>/dev/null | tail -1
