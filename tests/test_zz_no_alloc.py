"""GPU: dh_gemm_swiglu_streamk_noalloc (include/dualhyp.h) -- the stream-K SwiGLU decode GEMM with its scratch in a static device
array, for callers that may not allocate.  No counterpart in the reference, so the
checker is the existing entry point: dh_gemm with a caller-owned workspace runs the same kernel (bit-equal), dh_gemm without one
the cluster split-K kernel (same math, other summation order: the bf16 tolerance of test_gemm_streamk_swiglu_decode).

Collected last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

from dualhyp_b200 import _lib, ops, synthetic

pytestmark = pytest.mark.gpu
DEV = "cuda"
NOALLOC_SCRATCH_BYTES = 36 << 20          # DH_NOALLOC_SCRATCH_BYTES


def rnd(shape, stream, bound=1.0, seed=29):
    return synthetic.hash_uniform(shape, seed, stream, bound, device=DEV).to(torch.bfloat16)


def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def interleave(w1, w2):
    I = w1.shape[0]
    return torch.stack((w1.view(I // 32, 32, -1), w2.view(I // 32, 32, -1)), dim=1).reshape(2 * I, -1).contiguous()


def swiglu_args(x, W12, act, workspace=None):
    a = _lib.GemmArgs()
    a.A, a.lda = x.data_ptr(), x.stride(0)
    a.B, a.ldb = W12.data_ptr(), W12.stride(0)
    a.M, a.N, a.K1, a.K2 = x.shape[0], W12.shape[0], x.shape[1], 0
    a.epilogue, a.alpha = ops.EPI_SWIGLU, 1.0
    a.out2, a.ldo2 = act.data_ptr(), act.stride(0)
    if workspace is not None:
        a.workspace, a.workspace_bytes = workspace.data_ptr(), workspace.numel()
    return a


def call(fn_name, a):
    rc = getattr(_lib.load(), fn_name)(C.byref(a), torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, fn_name)


def run_pair(M, I, Cdim, stream):
    """One problem through the no-alloc entry point and through both forms of the existing one."""
    x = rnd((M, Cdim), stream)
    W1, W2 = rnd((I, Cdim), stream + 1, bound=1 / 32), rnd((I, Cdim), stream + 2, bound=1 / 32)
    W12 = interleave(W1, W2)
    got, same, plain = (torch.full((M, I), float("nan"), device=DEV, dtype=torch.bfloat16) for _ in range(3))
    ws = torch.zeros(ops.gemm_streamk_workspace_bytes(M, 2 * I, Cdim), device=DEV, dtype=torch.uint8)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    call("dh_gemm_swiglu_streamk_noalloc", swiglu_args(x, W12, got))
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    call("dh_gemm", swiglu_args(x, W12, same, ws))
    call("dh_gemm", swiglu_args(x, W12, plain))
    torch.cuda.synchronize()
    ref = F.silu(x.float() @ W1.float().t()) * (x.float() @ W2.float().t())
    return got, same, plain, ref, free_before - free_after


def test_noalloc_streamk_swiglu_matches_dh_gemm_on_two_calls_in_a_row():
    _lib.init(0)
    # TinyLlama's fc_1|fc_2 at 256 rows (33.0 MiB of the 36 MiB array), then a smaller problem of another tile count on the SAME
    # array (stale partial tiles and counters of the first call must not leak), then the first shape again with new inputs
    leaked = []
    for i, (M, I, Cdim) in enumerate([(256, 5632, 2048), (16, 1024, 512), (256, 5632, 2048), (100, 5632, 2048)]):
        got, same, plain, ref, d_free = run_pair(M, I, Cdim, 900 + 10 * i)
        assert not torch.isnan(got.float()).any(), f"call {i}: output rows left unwritten"
        assert torch.equal(got, same), f"call {i}: differs from dh_gemm with a caller workspace (same kernel, must be bit-equal)"
        assert rel(got, plain) < 8e-3, f"call {i}: rel {rel(got, plain):.3e} against the split-K kernel"
        assert rel(got, ref) < 8e-3, f"call {i}: rel {rel(got, ref):.3e} against fp32 math"
        leaked.append(d_free)
    # the first call on a device may load the module that holds the array; later calls must not move the free-memory counter
    assert all(d == 0 for d in leaked[1:]), f"device memory taken by a no-alloc call: {leaked}"


def test_noalloc_streamk_rejects_what_it_cannot_hold():
    _lib.init(0)
    lib = _lib.load()
    M, I, Cdim = 256, 147 * 64, 2048                         # 147 tiles * 3 slices * 2 row blocks * 64 KiB > 36 MiB
    assert ops.gemm_streamk_workspace_bytes(M, 2 * I, Cdim) > NOALLOC_SCRATCH_BYTES
    x, W12 = rnd((M, Cdim), 990), rnd((2 * I, Cdim), 991, bound=1 / 32)
    act = torch.zeros((M, I), device=DEV, dtype=torch.bfloat16)
    rc = lib.dh_gemm_swiglu_streamk_noalloc(C.byref(swiglu_args(x, W12, act)), torch.cuda.current_stream().cuda_stream)
    assert rc != 0 and "DH_NOALLOC_SCRATCH_BYTES" in _lib.last_error()
    a = swiglu_args(x[:, :512].contiguous(), W12[:1024, :512].contiguous(), act)
    a.epilogue = ops.EPI_BF16
    rc = lib.dh_gemm_swiglu_streamk_noalloc(C.byref(a), torch.cuda.current_stream().cuda_stream)
    assert rc != 0 and "DH_EPI_SWIGLU" in _lib.last_error()
    torch.cuda.synchronize()
    assert float(act.float().abs().sum()) == 0.0             # nothing was launched
