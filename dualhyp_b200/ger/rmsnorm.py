"""`ger.rmsnorm` mirror (reference ger/rmsnorm.py:4-24): same constructor, parameter name and formula,
computed by the vectorised sm_100a kernel (`dh_rmsnorm_fwd` / `dh_rmsnorm_bwd`)."""
import torch

from .. import ops


class _RMSNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, eps):
        C = x.shape[-1]
        x2 = x.reshape(-1, C).contiguous()
        if x2.dtype not in (torch.float32, torch.bfloat16):
            x2 = x2.float()
        w = weight.detach().float().contiguous()
        y = torch.empty(x2.shape, device=x2.device, dtype=torch.bfloat16)
        rstd = torch.empty(x2.shape[0], device=x2.device, dtype=torch.float32)
        ops.rmsnorm_fwd(x2, w, y, rstd, eps)
        ctx.save_for_backward(x2, w, rstd)
        ctx.in_shape, ctx.in_dtype = x.shape, x.dtype
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, w, rstd = ctx.saved_tensors
        dy2 = dy.reshape(x2.shape).to(torch.bfloat16).contiguous()
        dx = torch.empty(x2.shape, device=x2.device, dtype=torch.float32)
        ops.rmsnorm_bwd(x2, w, rstd, dy2, None, dx, None)
        return dx.view(ctx.in_shape).to(ctx.in_dtype), None, None


class RMSNorm(torch.nn.Module):
    """y = weight * x * rsqrt(mean(x^2, -1) + eps).  Inside GPT.forward the engine calls the same kernel
    directly; this standalone forward returns bf16 (what the following autocast linear consumes) and is
    differentiable w.r.t. x (the weight is frozen on the DualHyp path)."""

    def __init__(self, size: int, dim: int = -1, eps: float = 1e-5) -> None:
        super().__init__()
        if dim != -1:
            raise NotImplementedError("RMSNorm over a dimension other than the last is not on the DualHyp path")
        self.weight = torch.nn.Parameter(torch.ones(size))
        self.eps = eps
        self.dim = dim

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return _RMSNormFn.apply(x, self.weight, self.eps)

    def reset_parameters(self):
        torch.nn.init.ones_(self.weight)
