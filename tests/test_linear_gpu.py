"""GPU parity of the tcgen05 linear layer (C ABI b200_linear_f16) against a plain PyTorch fp32 reference of the same op.
Tolerance: fp32 accumulation + one fp16 rounding on both sides -> results agree to 1 fp16 ulp (rtol 2e-3) except for
summation-order effects, bounded by atol 2e-3 * sqrt(K)/32 on O(1) outputs."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(x, W, bias, epi, res):
    from chirrup_b200 import ops

    acc = x.float() @ W.float().t()
    h = lambda t: t.half().float()
    if epi == ops.EPI_NONE:
        return (acc + (bias.float() if bias is not None else 0)).half()
    if epi == ops.EPI_TANH:
        return torch.tanh(h(acc + (bias.float() if bias is not None else 0))).half()
    if epi == ops.EPI_SIGMOID:
        return torch.sigmoid(h(acc + (bias.float() if bias is not None else 0))).half()
    if epi == ops.EPI_RELU_SQ:
        return (torch.relu(h(acc)) ** 2).half()
    return (res.float() + h(acc)).half()


SHAPES = [
    (1, 128, 64), (3, 8, 8), (64, 256, 128), (128, 128, 4096), (129, 136, 72), (256, 4096, 4096), (256, 96, 4096),
    (256, 4096, 96), (256, 480, 2048), (200, 2048, 480), (256, 16384, 2048), (256, 2048, 8192), (700, 768, 768),
    (64, 65536, 768), (17, 1024, 3072),
]


@pytest.mark.parametrize("M,N,K", SHAPES)
def test_linear_plain(M, N, K):
    from chirrup_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(M * 31 + N * 7 + K)
    x = (torch.randn(M, K, generator=g, device="cuda")).half()
    W = (torch.randn(N, K, generator=g, device="cuda") * K ** -0.5).half()
    y = ops.linear(x, W)
    torch.cuda.synchronize()
    ref = _ref(x, W, None, ops.EPI_NONE, None)
    assert y.shape == (M, N)
    assert torch.allclose(y.float(), ref.float(), rtol=2e-3, atol=2e-3), (y.float() - ref.float()).abs().max().item()


@pytest.mark.parametrize("epi", [0, 1, 2, 3, 4])
@pytest.mark.parametrize("M,N,K", [(256, 512, 1024), (37, 264, 200), (256, 2048, 2048)])
def test_linear_epilogues(epi, M, N, K):
    from chirrup_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(epi * 100 + M)
    x = torch.randn(M, K, generator=g, device="cuda").half()
    W = (torch.randn(N, K, generator=g, device="cuda") * K ** -0.5).half()
    bias = torch.randn(N, generator=g, device="cuda").half() if epi in (0, 1, 2) else None
    res = torch.randn(M, N, generator=g, device="cuda").half() if epi == 4 else None
    y = ops.linear(x, W, bias, epi, res)
    torch.cuda.synchronize()
    ref = _ref(x, W, bias, epi, res)
    assert torch.allclose(y.float(), ref.float(), rtol=3e-3, atol=3e-3), (y.float() - ref.float()).abs().max().item()


@pytest.mark.parametrize("M", [256, 200, 129, 512, 64])
def test_linear_group_tmix_shapes(M):
    """The grouped launch of one TMix layer (r, k, v + the four LoRA down-projections: 7 problems, mixed N and epilogues) and of
    the four up-projections (K = 96..480), each against the fp32 reference; M > 128 runs on CTA pairs (cta_group::2)."""
    from chirrup_b200 import ops

    C = 1024
    g = torch.Generator(device="cuda").manual_seed(M)
    rn = lambda *s: torch.randn(*s, generator=g, device="cuda")
    dims = [(C, ops.EPI_NONE), (C, ops.EPI_NONE), (C, ops.EPI_NONE), (128, ops.EPI_TANH), (128, ops.EPI_NONE),
            (480, ops.EPI_SIGMOID), (96, ops.EPI_NONE)]
    xs = [rn(M, C).half() for _ in dims]
    Ws = [(rn(N, C) * C ** -0.5).half() for N, _ in dims]
    ys = ops.linear_group(xs, Ws, epilogues=[e for _, e in dims])
    torch.cuda.synchronize()
    for x, W, (N, e), y in zip(xs, Ws, dims, ys):
        ref = _ref(x, W, None, e, None)
        assert torch.allclose(y.float(), ref.float(), rtol=3e-3, atol=3e-3), (N, e, (y.float() - ref.float()).abs().max().item())
    ups = [(128, ops.EPI_NONE, True), (128, ops.EPI_SIGMOID, True), (480, ops.EPI_NONE, False), (96, ops.EPI_SIGMOID, True)]
    xs = [rn(M, K).half() for K, _, _ in ups]
    Ws = [(rn(C, K) * K ** -0.5).half() for K, _, _ in ups]
    bs = [rn(C).half() if b else None for _, _, b in ups]
    ys = ops.linear_group(xs, Ws, biases=bs, epilogues=[e for _, e, _ in ups])
    torch.cuda.synchronize()
    for x, W, b, (K, e, _), y in zip(xs, Ws, bs, ups, ys):
        ref = _ref(x, W, b, e, None)
        assert torch.allclose(y.float(), ref.float(), rtol=3e-3, atol=3e-3), (K, e, (y.float() - ref.float()).abs().max().item())


@pytest.mark.parametrize("M,N,K", [(256, 4096, 16384), (256, 512, 512), (255, 264, 4096), (1024, 1024, 1024),
                                   (512, 4096, 16384), (500, 4096, 4096), (768, 1024, 8192)])
def test_linear_pair_splitk_and_tails(M, N, K):
    """CTA-pair kernel: K split over the cluster (deterministic thread-major partials), N / M tails, several M blocks."""
    from chirrup_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(N + K)
    x = torch.randn(M, K, generator=g, device="cuda").half()
    W = (torch.randn(N, K, generator=g, device="cuda") * K ** -0.5).half()
    res = torch.randn(M, N, generator=g, device="cuda").half()
    a = ops.linear(x, W, None, ops.EPI_RESADD, res)
    b = ops.linear(x, W, None, ops.EPI_RESADD, res)
    torch.cuda.synchronize()
    ref = _ref(x, W, None, ops.EPI_RESADD, res)
    assert torch.equal(a, b)
    assert torch.allclose(a.float(), ref.float(), rtol=3e-3, atol=3e-3), (a.float() - ref.float()).abs().max().item()


def test_linear_inplace_residual_and_determinism():
    from chirrup_b200 import ops

    M, N, K = 256, 1024, 4096
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn(M, K, generator=g, device="cuda").half()
    W = (torch.randn(N, K, generator=g, device="cuda") * K ** -0.5).half()
    res = torch.randn(M, N, generator=g, device="cuda").half()
    ref = _ref(x, W, None, ops.EPI_RESADD, res)
    outs = []
    for _ in range(3):
        buf = res.clone()
        ops.linear(x, W, None, ops.EPI_RESADD, residual=buf, out=buf)
        outs.append(buf)
    torch.cuda.synchronize()
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[1], outs[2])   # split-K reduction order is fixed
    assert torch.allclose(outs[0].float(), ref.float(), rtol=3e-3, atol=3e-3)


def test_linear_strided_rows_and_empty():
    from chirrup_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(4)
    big = torch.randn(40, 3, 256, generator=g, device="cuda").half()
    x = big[:, 1, :]                       # row pitch 768
    W = (torch.randn(64, 256, generator=g, device="cuda") / 16).half()
    y = ops.linear(x, W)
    assert torch.allclose(y.float(), (x.float() @ W.float().t()), rtol=3e-3, atol=3e-3)
    assert ops.linear(x[:0], W).shape == (0, 64)


def test_spmv_matches_exact_product():
    """b200_spmv_forward vs the C oracle's exact product (oracle/wkv7_oracle.c: wkv7_oracle_spmv): fp32 accumulation in a
    fixed order vs double -> equal to 1 fp16 ulp; run-to-run bit-identical (the reference's fp16 atomics are not)."""
    from chirrup_b200 import ops
    from oracle import rwkv7_oracle as O

    D, C = 3072, 768
    g = torch.Generator().manual_seed(8)
    vec = torch.relu(torch.randn(D, generator=g) - 0.8) ** 2
    vec = vec.half()
    mat = (torch.randn(D, C, generator=g) * 0.05).half()
    want = O.spmv_cpu(vec, mat)
    a = ops.spmv_forward(vec.cuda(), mat.cuda())
    b = ops.spmv_forward(vec.cuda(), mat.cuda())
    torch.cuda.synchronize()
    assert torch.equal(a, b)
    assert torch.allclose(a.cpu().float(), want.float(), rtol=2e-3, atol=1e-4)
    z = ops.spmv_forward(torch.zeros(D, dtype=torch.float16, device="cuda"), mat.cuda())
    assert torch.count_nonzero(z) == 0


@pytest.mark.parametrize("D,C,density", [(3072, 768, 0.2), (8192, 2048, 0.5), (16384, 4096, 0.1), (16384, 4096, 0.9), (4160, 264, 0.5),
                                           (40, 64, 1.0)])
def test_spmv_vs_reference_kernel_and_exact_product(D, C, density):
    """The bsz-1 FFN value product (SURVEY a-6; 512 CTAs in clusters of 8 at C = 4096): against the exact fp32 product (one
    rounding; ours must be within an fp16 ulp of it), against the REFERENCE'S OWN kernel spvecmatmul_noindices
    (oracle/_ref, Albatross/cuda/rwkv7_state_fwd_fp16.cu:222-310; it accumulates in fp16 with atomics, so only to its own
    noise), and bit-identical run to run — odd shapes, all-dense and very sparse vectors included."""
    import os

    from chirrup_b200 import ops
    from oracle import rwkv7_oracle as O

    g = torch.Generator().manual_seed(D + C)
    vec = torch.randn(D, generator=g).abs() ** 2
    vec[torch.rand(D, generator=g) > density] = 0
    vec[D // 3] = -0.0                                                 # a negative zero is a zero (reference .cu:247-249)
    vec = vec.half().cuda()
    mat = (torch.randn(D, C, generator=g) * 0.05).half().cuda()
    exact = (vec.float() @ mat.float())
    a = ops.spmv_forward(vec, mat)
    b = ops.spmv_forward(vec, mat)
    torch.cuda.synchronize()
    assert torch.equal(a, b)
    scale = exact.abs().max().clamp_min(1e-6)
    assert ((a.float() - exact).abs().max() / scale) <= 1e-3
    assert ((a.float() - exact).abs() <= exact.abs() * 2e-3 + 2e-4 * scale).all()                    # fp32 sums in another order
    ref_so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libref_wkv7.so")
    if os.path.exists(ref_so) and C % 256 == 0 and D % 64 == 0:       # the reference asserts these (.cu:325-326)
        r = O.spmv_ref_cuda(vec, mat)
        err_ref = float((r.float() - exact).abs().max() / scale)       # the reference's own fp16-accumulation noise
        err_ours = float((a.float() - exact).abs().max() / scale)
        assert err_ours <= err_ref + 1e-3 and float((a.float() - r.float()).abs().max() / scale) <= 2 * err_ref + 1e-3


def test_linear_fuzz_random_shapes_and_epilogues():
    """60 seeded random problems across the three GEMM kernels (one CTA per tile for M <= 128, CTA pairs with K split for
    128 < M, persistent pairs for many M blocks): M, N, K with ragged tails (N, K multiples of 8 only), every epilogue; each
    result must match the fp32 torch reference and be bit-identical when the call is repeated."""
    import random

    from chirrup_b200 import ops

    rnd = random.Random(20240607)
    g = torch.Generator(device="cuda").manual_seed(77)
    for case in range(60):
        M = rnd.choice([rnd.randint(1, 128), rnd.randint(129, 256), rnd.randint(257, 700), rnd.randint(701, 2600)])
        N = 8 * rnd.randint(1, 96) if rnd.random() < 0.7 else 8 * rnd.randint(97, 600)
        K = 8 * rnd.randint(1, 80) if rnd.random() < 0.6 else 8 * rnd.randint(81, 640)
        epi = rnd.randint(0, 4)
        x = torch.randn(M, K, generator=g, device="cuda").half()
        W = (torch.randn(N, K, generator=g, device="cuda") * K ** -0.5).half()
        bias = torch.randn(N, generator=g, device="cuda").half() if epi in (0, 1, 2) and rnd.random() < 0.7 else None
        res = torch.randn(M, N, generator=g, device="cuda").half() if epi == 4 else None
        a = ops.linear(x, W, bias, epi, res)
        b = ops.linear(x, W, bias, epi, res)
        torch.cuda.synchronize()
        ref = _ref(x, W, bias, epi, res)
        assert torch.equal(a, b), (case, M, N, K, epi)
        err = (a.float() - ref.float()).abs().max().item()
        assert torch.allclose(a.float(), ref.float(), rtol=4e-3, atol=4e-3), (case, M, N, K, epi, err)
