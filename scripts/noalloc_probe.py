"""Device time of omr_exact_topk_noalloc on the largest scan it accepts (diagnostics)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from omnia_b200 import synth, _lib as L
from omnia_b200.store import exact_topk_noalloc
n, dim, k = L.NOALLOC_MAX_ROWS, int(os.environ.get("DIM", 384)), 10
rows = torch.from_numpy(synth.gauss_unit(n, dim)).cuda()
q = rows[12345].clone()
o_r = torch.empty(k, dtype=torch.int64, device="cuda"); o_d = torch.empty(k, dtype=torch.float64, device="cuda")
o_c = torch.empty(1, dtype=torch.int32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    exact_topk_noalloc(rows, q, k, o_r, o_d, o_c, st)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    exact_topk_noalloc(rows, q, k, o_r, o_d, o_c, st)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
print(json.dumps({"entry": "omr_exact_topk_noalloc", "rows": n, "dim": dim, "k": k, "ms_per_call": ms,
                  "gb_per_s": n * dim * 4 / ms / 1e6, "first_row": int(o_r[0])}))
