#!/bin/bash
# knock-out timing of the cluster GEMM: which stage costs what (results are garbage, only the time counts)
V=$PWD/video_flow_ml_b200/variants
for shape in "42240 486 324" "21120 768 512"; do
for v in base NOSTORE NOEPI NOEPI_NOSTORE; do
  for cs in 4 8; do
    if [ $v = base ]; then L=$PWD/video_flow_ml_b200/libvfml.so; else L=$V/libvfml_ws_$v.so; fi
    echo "== $v CS=$cs"
    VFML_WS_CS=$cs VFML_LIB_PATH=$L VFML_GEMM_TRACE=1 python - $shape <<'PY' 2>&1 | grep "trace\]" | tail -3 | cut -c1-330
import sys, torch
sys.path.insert(0, ".")
from video_flow_ml_b200 import ops
M, N, K = (int(a) for a in sys.argv[1:4])
buf = lambda n, c: torch.randn((n, (c + 7) // 8 * 8), device="cuda").half()[:, :c]
x, w, o, r = buf(M, K), buf(N, K), buf(M, N), buf(M, N)
b = torch.randn(N, device="cuda")
for _ in range(2): ops.gemm(x, w, o)
ops.gemm(x, w, o)
ops.gemm(x, w, o, bias=b, act="gelu")
ops.gemm(x, w, o, bias=b, res=r, act="gelu")
torch.cuda.synchronize()
PY
  done
done
done
