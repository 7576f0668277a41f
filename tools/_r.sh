cat > /tmp/fe_t.py <<'PY'
import torch, numpy as np, sys
from mrs_optic_flow_b200 import api
dev = torch.device("cuda", 0)
for n in (64, 512, 2048):
    raw = torch.randint(0, 256, (n, 480, 752, 3), dtype=torch.uint8, device=dev)
    be = api.Backend(480, 120, 80.0)
    for _ in range(3): be.frontend(raw, 1, 376)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): be.frontend(raw, 1, 376)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(n, "frames:", round(ms * 1e3, 1), "us,", round(n * 480 * 480 * 4 / (ms * 1e-3) / 1e9), "GB/s")
    be.close(); del raw
PY
echo "A (words):"; timeout 200 python /tmp/fe_t.py
echo "B (bytes):"; MOF_B200_LIB=$PWD/ab_variant.so timeout 200 python /tmp/fe_t.py
