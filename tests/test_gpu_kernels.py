"""Every hand-written kernel against a plain NumPy / PyTorch restatement of the same op."""
import numpy as np
import pytest
import torch

from cme213_sp18_b200.ops import host
from cme213_sp18_b200.utils import heat as heat_utils

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    from cme213_sp18_b200.ops import device

    return device


def cuda(a):
    return torch.from_numpy(np.array(a, copy=True)).cuda()


# ------------------------------------------------------------------ hw2 q1 shift
@pytest.mark.parametrize("variant", ["char", "uint", "uint2", "uint4"])
@pytest.mark.parametrize("n", [1, 15, 16, 4099, 1_235_150])
def test_shift_exact(dev, variant, n):
    rng = np.random.default_rng(n)
    data = rng.integers(0, 256, size=n, dtype=np.uint8)      # includes bytes that overflow (Q13)
    for shift in (1, 25, 200):
        out = dev.shift_bytes(cuda(data), shift, variant).cpu().numpy()
        assert np.array_equal(out, host.shift_bytes(data, shift))


# ------------------------------------------------------------------ hw2 q2 pagerank
@pytest.mark.parametrize("n,avg", [(1 << 12, 2), (1 << 15, 8), (50_000, 19)])
def test_pagerank_100_ulp(dev, n, avg):
    from cme213_sp18_b200.models.pagerank import pagerank_device

    rp, edges, inv = host.pagerank_graph(n, avg, seed=3)
    r0 = np.full(n, 1.0 / n, dtype=np.float32)
    want = host.pagerank(rp, edges, r0, inv, 6)
    got = pagerank_device(cuda(rp), cuda(edges.view(np.int32)), cuda(r0), cuda(inv), 6).cpu().numpy()
    ulp, where = host.max_ulp(got, want)
    assert ulp <= 100, (ulp, where)     # ref hw2code/main_q2.cu:28 kMaxUlps


def test_pagerank_device_graph_matches_host(dev):
    n, avg, seed = 10_000, 5, 42
    rp, edges, _ = host.pagerank_graph(n, avg, seed=seed)
    e = torch.empty(len(edges), dtype=torch.int32, device="cuda")
    dev.pagerank_gen_edges(e, 0, n, seed)
    assert np.array_equal(e.cpu().numpy().view(np.uint32), edges)


# ------------------------------------------------------------------ hw3 stencil
@pytest.mark.parametrize("order", [2, 4, 8])
@pytest.mark.parametrize("variant", ["global", "loop", "shared"])
@pytest.mark.parametrize("shape", [(256, 256), (300, 77), (1000, 130)])
def test_heat_512_ulp(dev, order, variant, shape):
    from cme213_sp18_b200.models.heat import heat_device

    nx, ny = shape
    p = heat_utils.SimParams(nx=nx, ny=ny, iters=40, order=order)
    g0 = host.heat_init(p.gx, p.gy, p.dx, p.dy)
    want = host.heat_cpu(g0, order, p.iters, p.xcfl, p.ycfl, p.bc_scale, "gpu")
    for fused_bc in (False, True):
        got = heat_device(cuda(g0), p, variant=variant, fused_bc=fused_bc).cpu().numpy()
        ulp, where = host.max_ulp(got, want)
        assert ulp <= 512, (ulp, where, fused_bc)   # ref hw3code/main.cu:217-266


# ------------------------------------------------------------------ hw4 cipher
def test_sanitize_and_histogram(dev, english_text):
    raw = np.frombuffer(english_text, dtype=np.uint8)
    lower = np.where((raw >= 65) & (raw <= 90), raw + 32, raw)
    clean = lower[(lower >= 97) & (lower <= 122)]
    buf, count = dev.sanitize_compact(cuda(raw))
    n = int(count.item())
    assert n == len(clean)
    assert np.array_equal(buf[:n].cpu().numpy(), clean)
    h = dev.byte_histogram(buf[:n]).cpu().numpy()
    assert np.array_equal(h, np.bincount(clean, minlength=256))
    for period in (4, 7, 13):
        hs = dev.strided_histograms(buf[:n], period).cpu().numpy()
        for k in range(period):
            assert np.array_equal(hs[k], np.bincount(clean[k::period], minlength=256))
        assert np.array_equal(dev.byte_histogram(buf[:n], stride=period, offset=2).cpu().numpy(), hs[2])


def test_vigenere_and_lags(dev, english_text):
    raw = np.frombuffer(english_text, dtype=np.uint8)
    clean = raw[(raw >= 97) & (raw <= 122)][:100_001]
    shifts = np.array([3, 25, 1, 17, 9], dtype=np.int32)
    enc = dev.vigenere_shift(cuda(clean), cuda(shifts)).cpu().numpy()
    want = (clean.astype(np.int32) + shifts[np.arange(len(clean)) % 5]).astype(np.uint8)
    assert np.array_equal(enc, want)
    dec = dev.vigenere_shift(cuda(enc), cuda(-shifts)).cpu().numpy()
    assert np.array_equal(dec, clean)
    w = dev.vigenere_shift(cuda(clean), cuda(shifts), wrap26=True).cpu().numpy()
    assert w.min() >= 97 and w.max() <= 122
    lags = dev.lag_match_counts(cuda(enc), 4, 20).cpu().numpy()
    for k in range(20):
        lag = 4 + k
        assert lags[k] == int((enc[:-lag] == enc[lag:]).sum())


# ------------------------------------------------------------------ hw1 on device
def test_parity_and_radix(dev):
    rng = np.random.default_rng(1)
    v = rng.integers(0, 101, size=3_000_001, dtype=np.uint32)
    got = dev.parity_sum(cuda(v.view(np.int32))).cpu().numpy()
    assert (int(got[0]), int(got[1])) == host.parity_sum(v)[0]
    for n in (1, 255, 4096, 4097, 1_000_003):
        k = rng.integers(0, 2**32, size=n, dtype=np.uint64).astype(np.uint32)
        out = dev.radix_sort(cuda(k.view(np.int32))).cpu().numpy().view(np.uint32)
        assert np.array_equal(out, np.sort(k))


# ------------------------------------------------------------------ fp64 GEMM (ref -g 4)
@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
@pytest.mark.parametrize("mnk", [(800, 1000, 784), (800, 10, 1000), (37, 65, 19), (1, 1, 1),
                                 (10, 1000, 800),       # wide C: the skinny kernel on the transposed product
                                 (2100, 2000, 70),      # enough 128 x 128 tiles to fill the machine
                                 (16, 16, 300), (17, 9, 5), (1000, 100, 10)])
def test_dgemm_rel_inf_norm(dev, ta, tb, mnk):
    M, N, K = mnk
    g = torch.Generator().manual_seed(M * 7 + N)
    A = torch.randn((M, K) if not ta else (K, M), dtype=torch.float64, generator=g)
    B = torch.randn((K, N) if not tb else (N, K), dtype=torch.float64, generator=g)
    C0 = torch.randn((M, N), dtype=torch.float64, generator=g)
    alpha, beta = 2.0, 5.0                                   # ref fpcode/utils/tests.cpp:131-132
    want = alpha * ((A.T if ta else A) @ (B.T if tb else B)) + beta * C0
    # column-major device buffers: a torch (r, c) matrix in column-major is its transpose, contiguous
    colmajor = lambda t: t.T.contiguous().cuda()             # noqa: E731
    dA, dB, dC = colmajor(A), colmajor(B), colmajor(C0)
    dev.dgemm(ta, tb, M, N, K, alpha, dA, A.shape[0], dB, B.shape[0], beta, dC, M)
    got = dC.cpu().T
    rel = (got - want).abs().sum(dim=1).max() / want.abs().sum(dim=1).max()
    assert rel <= 1e-12, float(rel)                          # ref GEMM_TOL
    # every element, not only the norm: a misplaced fragment would hide in a norm of random data
    scale = alpha * (A.abs().T if ta else A.abs()) @ (B.abs().T if tb else B.abs()) + beta * C0.abs()
    assert ((got - want).abs() <= 1e-13 * scale + 1e-300).all()
    # the k-split of the skinny kernel adds in a fixed order: a second call gives the same bits
    dC2 = colmajor(C0)
    dev.dgemm(ta, tb, M, N, K, alpha, dA, A.shape[0], dB, B.shape[0], beta, dC2, M)
    assert torch.equal(dC, dC2)


def test_dgemm_beta_zero_ignores_nan(dev):
    A = torch.randn(8, 8, dtype=torch.float64).cuda()
    C = torch.full((8, 8), float("nan"), dtype=torch.float64).cuda()
    dev.dgemm(False, False, 8, 8, 8, 1.0, A, 8, A, 8, 0.0, C, 8)
    assert torch.isfinite(C).all()


# ------------------------------------------------------------------ tcgen05 GEMMs
@pytest.mark.parametrize("mnk", [(128, 64, 64), (128, 256, 128), (300, 128, 784), (1000, 320, 100),
                                 (4096, 1024, 785), (129, 64, 17)])
def test_gemm_bf16_kmajor(dev, mnk):
    M, N, K = mnk
    g = torch.Generator().manual_seed(M + N + K)
    ldk = (K + 63) // 64 * 64
    A = torch.zeros((M, ldk), dtype=torch.bfloat16)
    B = torch.zeros((N, ldk), dtype=torch.bfloat16)
    A[:, :K] = torch.randn((M, K), generator=g).to(torch.bfloat16)
    B[:, :K] = torch.randn((N, K), generator=g).to(torch.bfloat16)
    # poison the pad columns: the kernel must bound K by the tensor map, not by ld
    A[:, K:] = 1e3
    B[:, K:] = 1e3
    dA, dB = A.cuda(), B.cuda()
    out = dev.gemm_bf16(dA[:, :K], dB[:, :K])
    want = A[:, :K].float() @ B[:, :K].float().T
    err = (out.float().cpu() - want).abs().max() / want.abs().max()
    assert err < 1e-2, float(err)


def test_mlp_fwd1_sigmoid(dev):
    rows, K, H = 777, 785, 100
    ldx, ldh = 832, 128
    g = torch.Generator().manual_seed(0)
    X = torch.zeros((rows, ldx), dtype=torch.bfloat16)
    X[:, :K - 1] = torch.randint(0, 256, (rows, K - 1), generator=g).to(torch.bfloat16)
    X[:, K - 1] = 1.0
    W = torch.zeros((128, ldx), dtype=torch.bfloat16)
    W[:H, :K] = (0.01 * torch.randn((H, K), generator=g)).to(torch.bfloat16)
    A1 = torch.zeros((rows, ldh), dtype=torch.bfloat16, device="cuda")
    dev.mlp_fwd1(X.cuda(), W.cuda(), A1, rows, K, H)
    want = torch.sigmoid(X[:, :K].float() @ W[:H, :K].float().T)
    got = A1[:, :H].float().cpu()
    assert (got - want).abs().max() < 1e-2


@pytest.mark.parametrize("rows,K,H", [(64, 64, 128), (1000, 785, 100), (5000, 785, 1024), (333, 100, 64)])
def test_mlp_dw1_streamk(dev, rows, K, H):
    g = torch.Generator().manual_seed(rows)
    ldx = (K + 63) // 64 * 64
    ldh = (H + 63) // 64 * 64
    X = torch.zeros((rows, ldx), dtype=torch.bfloat16)
    X[:, :K] = torch.randn((rows, K), generator=g).to(torch.bfloat16)
    dZ = torch.zeros((rows, ldh), dtype=torch.bfloat16)
    dZ[:, :H] = torch.randn((rows, H), generator=g).to(torch.bfloat16)
    dZ[:, H:] = 7.0                                          # pad columns must not leak
    G = torch.zeros((H, ldx), dtype=torch.float32, device="cuda")
    dev.mlp_dw1(dZ.cuda(), X.cuda(), G, rows, K, H, ldx)
    want = dZ[:, :H].float().T @ X[:, :K].float()
    got = G[:, :K].cpu()
    err = (got - want).abs().max() / want.abs().max()
    assert err < 1e-3, float(err)
    if ldx > K:
        assert float(G[:, K:].abs().max()) == 0.0


def test_u8_to_bf16_rows(dev):
    rows, cols, ld = 37, 784, 832
    src = torch.randint(0, 256, (rows, cols), dtype=torch.uint8, device="cuda")
    dst = torch.full((rows, ld), 9.0, dtype=torch.bfloat16, device="cuda")
    dev.u8_to_bf16_rows(src.data_ptr(), dst, rows, cols, add_ones=True)
    assert torch.equal(dst[:, :cols].float(), src.float())
    assert float(dst[:, cols].float().min()) == 1.0 and float(dst[:, cols + 1:].float().abs().max()) == 0.0
    # odd column count takes the scalar path
    src2 = torch.randint(0, 256, (5, 30), dtype=torch.uint8, device="cuda")
    dst2 = torch.full((5, 64), 9.0, dtype=torch.bfloat16, device="cuda")
    dev.u8_to_bf16_rows(src2.data_ptr(), dst2, 5, 30, add_ones=False)
    assert torch.equal(dst2[:, :30].float(), src2.float()) and float(dst2[:, 30:].float().abs().max()) == 0.0


@pytest.mark.parametrize("rows,H", [(50, 100), (1000, 1024), (257, 2500)])
def test_fwd2_softmax_and_bwd_elem(dev, rows, H):
    C = 10
    ldh = (H + 63) // 64 * 64
    g = torch.Generator().manual_seed(H)
    A1 = torch.full((rows, ldh), 0.5, dtype=torch.bfloat16)      # pad columns hold sigmoid(0)
    A1[:, :H] = torch.rand((rows, H), generator=g).to(torch.bfloat16)
    W2 = (0.1 * torch.randn((C, H), generator=g)).float()
    b2 = torch.randn(C, generator=g).float()
    labels = torch.randint(0, C, (rows,), generator=g).int()
    scale = 1.0 / rows
    dz2 = torch.zeros((rows, 16), device="cuda")
    loss = torch.zeros(1, device="cuda")
    pred = torch.zeros(rows, dtype=torch.int32, device="cuda")
    gb2 = torch.zeros(C, device="cuda")
    dA1, dW2 = A1.cuda(), W2.cuda()
    dev.mlp_fwd2_softmax(dA1, dW2, b2.cuda(), labels.cuda(), dz2, loss, pred, gb2, scale, rows, H)
    z = A1[:, :H].float() @ W2.T + b2
    p = torch.softmax(z, dim=1)
    want_dz2 = (p - torch.nn.functional.one_hot(labels.long(), C)) * scale
    assert (dz2[:, :C].cpu() - want_dz2).abs().max() < 1e-5
    assert float(dz2[:, C:].abs().max()) == 0.0
    want_loss = -torch.log(p[torch.arange(rows), labels.long()]).sum()
    assert abs(float(loss.item()) - float(want_loss)) < 1e-3 * max(1.0, float(want_loss))
    assert torch.equal(pred.cpu().long(), z.argmax(dim=1))
    assert (gb2.cpu() - want_dz2.sum(dim=0)).abs().max() < 1e-5

    dZ1 = torch.zeros((rows, ldh), dtype=torch.bfloat16, device="cuda")
    gW2 = torch.zeros((C, H), device="cuda")
    dev.mlp_bwd_elem(dA1, dz2, dW2, dZ1, gW2, rows, H)
    a = A1[:, :H].float()
    want_dZ1 = (want_dz2 @ W2) * a * (1 - a)
    got_dZ1 = dZ1[:, :H].float().cpu()
    assert (got_dZ1 - want_dZ1).abs().max() <= 1e-2 * want_dZ1.abs().max() + 1e-9
    want_gW2 = want_dz2.T @ a
    assert (gW2.cpu() - want_gW2).abs().max() < 1e-4 * max(1.0, float(want_gW2.abs().max()))


@pytest.mark.parametrize("rows,H", [(50, 100), (1000, 1024), (700, 2500), (4096, 8192)])
def test_tensor_core_tail_fwd2_and_dw2(dev, rows, H):
    """Layer 2 + softmax epilogue and the dW2 product on tcgen05 vs fp32 PyTorch."""
    C = 10
    ldh = (H + 63) // 64 * 64
    g = torch.Generator().manual_seed(H + rows)
    A1 = torch.full((rows, ldh), 0.5, dtype=torch.bfloat16)      # pad columns hold sigmoid(0)
    A1[:, :H] = torch.rand((rows, H), generator=g).to(torch.bfloat16)
    W2 = (0.1 * torch.randn((C, H), generator=g)).float()
    b2 = torch.randn(C, generator=g).float()
    labels = torch.randint(0, C, (rows,), generator=g).int()
    scale = 1.0 / rows
    dW2 = W2.cuda()
    w2c = torch.zeros((16, ldh), dtype=torch.bfloat16, device="cuda")
    dev.f32_to_bf16_2d(dW2, H, w2c, ldh, C, H)
    assert torch.equal(w2c[:C, :H].float().cpu(), W2.bfloat16().float())
    assert float(w2c[C:].float().abs().max()) == 0.0
    dz2 = torch.zeros((rows, 16), device="cuda")
    dz2b = torch.zeros((rows, 64), dtype=torch.bfloat16, device="cuda")
    loss = torch.zeros(1, device="cuda")
    pred = torch.zeros(rows, dtype=torch.int32, device="cuda")
    gb2 = torch.zeros(C, device="cuda")
    dA1 = A1.cuda()
    dev.mlp_fwd2_tc(dA1, w2c, b2.cuda(), labels.cuda(), dz2, dz2b, loss, pred, gb2, scale, rows, H)
    z = A1[:, :H].float() @ W2.bfloat16().float().T + b2
    p = torch.softmax(z, dim=1)
    want_dz2 = (p - torch.nn.functional.one_hot(labels.long(), C)) * scale
    assert (dz2[:, :C].cpu() - want_dz2).abs().max() < 2e-3 * scale + 1e-7
    assert float(dz2[:, C:].abs().max()) == 0.0
    assert torch.equal(dz2b[:, :16].float().cpu(), dz2.cpu().bfloat16().float())
    want_loss = -torch.log(p[torch.arange(rows), labels.long()]).sum()
    assert abs(float(loss.item()) - float(want_loss)) < 2e-3 * max(1.0, float(want_loss))
    agree = (pred.cpu().long() == z.argmax(dim=1)).double().mean()
    assert agree > 0.995
    assert (gb2.cpu() - want_dz2.sum(dim=0)).abs().max() < 1e-4

    gW2 = torch.zeros((C, H), device="cuda")
    dev.mlp_dw2_tc(dA1, dz2b, gW2, rows, H)
    want_gW2 = dz2b[:, :C].float().cpu().T @ A1[:, :H].float()
    err = (gW2.cpu() - want_gW2).abs().max() / want_gW2.abs().max()
    assert err < 1e-3, float(err)
    # dZ1-only backward pass
    dZ1 = torch.zeros((rows, ldh), dtype=torch.bfloat16, device="cuda")
    dev.mlp_bwd_elem(dA1, dz2, dW2, dZ1, None, rows, H)
    a = A1[:, :H].float()
    want_dZ1 = (dz2[:, :C].cpu() @ W2) * a * (1 - a)
    assert (dZ1[:, :H].float().cpu() - want_dZ1).abs().max() <= 1e-2 * want_dZ1.abs().max() + 1e-9
    # the same pass on the tensor cores
    w2t = torch.zeros((ldh, 64), dtype=torch.bfloat16, device="cuda")
    dev.mlp_recast_w2(dW2, w2c, w2t, H)
    assert torch.equal(w2t[:H, :C].float().cpu(), W2.T.bfloat16().float())
    dZ1b = torch.full((rows, ldh), 3.0, dtype=torch.bfloat16, device="cuda")
    dev.mlp_bwd_dz1_tc(dA1, dz2b, w2t, dZ1b, rows, H)
    want_tc = (dz2b[:, :C].float().cpu() @ W2.bfloat16().float()) * a * (1 - a)
    assert (dZ1b[:, :H].float().cpu() - want_tc).abs().max() <= 1e-2 * want_tc.abs().max() + 1e-9
    h8 = (H + 7) // 8 * 8        # TMA stores clip the inner dimension at 16-byte granularity
    if ldh > h8:
        assert float((dZ1b[:, h8:].float() - 3.0).abs().max()) == 0.0   # pad columns are not written
    # dZ1 and dW2 fused into one pass
    dZ1f = torch.full((rows, ldh), 3.0, dtype=torch.bfloat16, device="cuda")
    gW2f = torch.zeros((C, H), device="cuda")
    dev.mlp_bwd_tail_tc(dA1, dz2b, w2t, dZ1f, gW2f, H, rows, H)
    assert torch.equal(dZ1f[:, :H], dZ1b[:, :H])
    errf = (gW2f.cpu() - want_gW2).abs().max() / want_gW2.abs().max()
    assert errf < 1e-3, float(errf)


@pytest.mark.parametrize("rows,H", [(256, 256), (1000, 1024), (4096, 2048), (8192, 8192), (777, 1000)])
def test_fwd_fused_rider_matches_two_kernel_forward(dev, rows, H):
    """A1 and the layer-2 outputs of the one-kernel forward (rider MMAs on the staged A1 tiles +
    per-row tail) against fp32 PyTorch, and bit-reproducible from call to call."""
    C, K = 10, 785
    ldx, ldh = 832, (H + 63) // 64 * 64
    g = torch.Generator().manual_seed(rows + H)
    X = torch.zeros((rows, ldx), dtype=torch.bfloat16)
    X[:, :K - 1] = torch.randint(0, 2, (rows, K - 1), generator=g).to(torch.bfloat16)
    X[:, K - 1] = 1.0
    W1 = torch.zeros((ldh, ldx), dtype=torch.bfloat16)
    W1[:H, :K] = (0.05 * torch.randn((H, K), generator=g)).to(torch.bfloat16)
    W2 = (0.1 * torch.randn((C, H), generator=g)).float()
    b2 = torch.randn(C, generator=g).float()
    labels = torch.randint(0, C, (rows,), generator=g).int()
    scale = 1.0 / rows
    w2c = torch.zeros((16, ldh), dtype=torch.bfloat16, device="cuda")
    dev.f32_to_bf16_2d(W2.cuda(), H, w2c, ldh, C, H)
    scratch = torch.zeros(dev.fwd_scratch_floats(rows), device="cuda")
    outs = []
    for _ in range(2):
        A1 = torch.full((rows, ldh), 7.0, dtype=torch.bfloat16, device="cuda")
        dz2 = torch.zeros((rows, 16), device="cuda")
        dz2b = torch.zeros((rows, 64), dtype=torch.bfloat16, device="cuda")
        loss = torch.zeros(1, device="cuda")
        pred = torch.zeros(rows, dtype=torch.int32, device="cuda")
        gb2 = torch.full((C,), 5.0, device="cuda")
        dev.mlp_fwd_fused(X.cuda(), W1.cuda(), w2c, b2.cuda(), labels.cuda(), A1, dz2, dz2b, loss, pred, gb2,
                          scratch, rows, scale, rows, K, H)
        torch.cuda.synchronize()
        outs.append((A1, dz2, dz2b, loss, pred, gb2))
    for a, b in zip(*outs):
        assert torch.equal(a, b)
    A1, dz2, dz2b, loss, pred, gb2 = outs[0]
    want_a1 = torch.sigmoid(X[:, :K].float() @ W1[:H, :K].float().T)
    assert (A1[:, :H].float().cpu() - want_a1).abs().max() < 1e-2
    a1 = A1[:, :H].float().cpu()
    z = a1 @ W2.bfloat16().float().T + b2
    p = torch.softmax(z, dim=1)
    want_dz2 = (p - torch.nn.functional.one_hot(labels.long(), C)) * scale
    assert (dz2[:, :C].cpu() - want_dz2).abs().max() < 2e-3 * scale + 1e-7
    assert float(dz2[:, C:].abs().max()) == 0.0
    assert torch.equal(dz2b[:, :16].float().cpu(), dz2.cpu().bfloat16().float())
    want_loss = -torch.log(p[torch.arange(rows), labels.long()]).sum()
    assert abs(float(loss.item()) - float(want_loss)) < 2e-3 * max(1.0, float(want_loss))
    assert (pred.cpu().long() == z.argmax(dim=1)).double().mean() > 0.995
    assert (gb2.cpu() - want_dz2.sum(dim=0)).abs().max() < 1e-4
