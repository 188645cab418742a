"""b200_linear_f16_static — the linear layer with its split-K scratch in a file-scope __device__ array (no allocation) —
against the existing entry point b200_linear_f16, through the C ABI, two calls in a row.  Collected last on purpose.
Tolerance: the one tests/test_linear_gpu.py uses for two evaluations of the same product (rtol/atol 3e-3: the two entry points
may pick different split-K plans, i.e. different summation orders); both are also checked against the fp32 PyTorch product."""
import pytest
import torch

pytestmark = pytest.mark.gpu

# (M, N, K, epilogue): split-K over a cluster of one-CTA tiles (M <= 128), CTA-pair plans with and without a K split, the
# persistent kernel (M > 256), a residual-add epilogue
CASES = [(64, 4096, 16384, 0), (17, 1024, 3072, 0), (256, 4096, 4096, 0), (256, 4096, 16384, 4), (256, 16384, 4096, 3),
         (700, 768, 768, 0), (128, 128, 4096, 2)]


def _call_static(L, x, W, out, bias, res, epi):
    return L.b200_linear_f16_static(x.shape[0], W.shape[0], x.shape[1], x.data_ptr(), x.stride(0), W.data_ptr(), W.stride(0),
                                    out.data_ptr(), out.stride(0), bias.data_ptr() if bias is not None else 0,
                                    res.data_ptr() if res is not None else 0, epi, torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("M,N,K,epi", CASES)
def test_static_scratch_linear_matches_workspace_linear_twice(M, N, K, epi):
    from chirrup_b200 import _lib, ops

    L = _lib.lib()
    assert L.b200_linear_static_scratch_bytes() == 32 * 1024 * 1024
    g = torch.Generator(device="cuda").manual_seed(M * 31 + N * 7 + K + epi)
    for call in range(2):   # the second call finds the first one's partials in the shared array
        x = torch.randn(M, K, generator=g, device="cuda").half()
        W = (torch.randn(N, K, generator=g, device="cuda") * K ** -0.5).half()
        bias = torch.randn(N, generator=g, device="cuda").half() if epi in (0, 1, 2) else None
        res = torch.randn(M, N, generator=g, device="cuda").half() if epi == 4 else None
        want = ops.linear(x, W, bias, epi, res)                      # b200_linear_f16 with a caller-provided workspace
        out = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        rc = _call_static(L, x, W, out, bias, res, epi)
        torch.cuda.synchronize()
        assert rc == 0, rc
        if call == 1:   # (the first call may load kernels / the module that holds the array: CUDA lazy loading)
            assert torch.cuda.mem_get_info()[0] == free0             # the call took no device memory
        assert torch.isfinite(out.float()).all()
        assert torch.allclose(out.float(), want.float(), rtol=3e-3, atol=3e-3), (call, (out.float() - want.float()).abs().max().item())
        acc = x.float() @ W.float().t()
        if epi == 0:
            ref = acc + bias.float()
        elif epi == 2:
            ref = torch.sigmoid((acc + bias.float()).half().float())
        elif epi == 3:
            ref = torch.relu(acc.half().float()) ** 2
        else:
            ref = res.float() + acc.half().float()
        assert torch.allclose(out.float(), ref.half().float(), rtol=3e-3, atol=3e-3), (call, (out.float() - ref).abs().max().item())


def test_static_scratch_linear_rejects_sizes_beyond_its_maximum():
    from chirrup_b200 import _lib

    L = _lib.lib()
    x = torch.zeros(8, 64, dtype=torch.float16, device="cuda")
    too_large = -5   # B200_ERR_TOO_LARGE
    s = torch.cuda.current_stream().cuda_stream
    # the sizes are checked before any pointer is used
    assert L.b200_linear_f16_static(4097, 64, 64, x.data_ptr(), 64, x.data_ptr(), 64, x.data_ptr(), 64, 0, 0, 0, s) == too_large
    assert L.b200_linear_f16_static(8, 65544, 64, x.data_ptr(), 64, x.data_ptr(), 64, x.data_ptr(), 65544, 0, 0, 0, s) == too_large
    assert L.b200_linear_f16_static(8, 64, 16392, x.data_ptr(), 16392, x.data_ptr(), 16392, x.data_ptr(), 64, 0, 0, 0, s) == too_large
    torch.cuda.synchronize()
