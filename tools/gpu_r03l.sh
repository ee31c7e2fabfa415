#!/bin/bash
# r03l (1 GPU, one minute): the host-side I/O changes inside GPU processes -- extraction tests, an epoch through Trainer.train, the ark leg
out=gpurun_out; tag=r03l
timeout 120 python -m pytest tests/test_gpu_dropin.py tests/test_gpu_train_api.py -x -q -m gpu -k "extract or epoch or cli" > $out/${tag}_pytest.log 2>&1; tail -2 $out/${tag}_pytest.log | cut -c1-300
timeout 60 python - <<'PY' 2>/dev/null | tail -1
import json, os, sys
sys.path.insert(0, os.getcwd())
import bench
class A: pass
r = bench.leg_ark_e2e(A(), [0])
print(json.dumps({"fm": round(r["fm"]["frames_per_s"] / 1e6, 1), "cm": round(r["cm"]["frames_per_s"] / 1e6, 1), "reader_fm": round(r["fm"]["reader_alone_frames_per_s"] / 1e6, 1), "reader_cm": round(r["cm"]["reader_alone_frames_per_s"] / 1e6, 1)}))
PY
