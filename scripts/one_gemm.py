"""Launch one decoder GEMM (+ fused epilogue) a few times: the target for `ncu -k regex:gemm_pair`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dualhyp_b200 import ops  # noqa: E402

DEV, bf = "cuda", torch.bfloat16
which = sys.argv[1] if len(sys.argv) > 1 else "swiglu_bwd"
M, C, I = 8192, 2048, 5632
n = torch.randn(M, C, device=DEV, dtype=bf)
if which == "swiglu_bwd":
    h12 = torch.randn(M, 2 * I, device=DEV, dtype=bf)
    w = (torch.randn(I, C, device=DEV, dtype=bf) * 0.02)
    out = torch.empty(M, 2 * I, device=DEV, dtype=bf)
    fn = lambda: ops.gemm(n, w, out, epilogue=ops.EPI_SWIGLU_BWD, aux=h12)
elif which == "swiglu":
    w = (torch.randn(2 * I, C, device=DEV, dtype=bf) * 0.02)
    h12 = torch.empty(M, 2 * I, device=DEV, dtype=bf)
    act = torch.empty(M, I, device=DEV, dtype=bf)
    fn = lambda: ops.gemm(n, w, h12, epilogue=ops.EPI_SWIGLU, out2=act)
elif which == "resid":
    w = (torch.randn(C, C, device=DEV, dtype=bf) * 0.02)
    x = torch.randn(M, C, device=DEV)
    out = torch.empty(M, C, device=DEV)
    fn = lambda: ops.gemm(n, w, out, epilogue=ops.EPI_RESID_F32, aux=x)
else:
    w = (torch.randn(2 * I, C, device=DEV, dtype=bf) * 0.02)
    out = torch.empty(M, 2 * I, device=DEV, dtype=bf)
    fn = lambda: ops.gemm(n, w, out)
for _ in range(3):
    fn()
torch.cuda.synchronize()
print("ok", which)
