"""GPU unit tests: every libexg.so kernel (through the C ABI, via exoground_b200.ops.backend) against
the torch restatement of its op spec (tests/fake_backend.py) on the same seeded inputs.
Tolerances: fp32 paths rel 1e-4 (north star); bf16 paths 2e-2 relative to the output scale."""
import math

import pytest
import torch

from exoground_b200 import ops
from exoground_b200.ops import lib as L
from tests.fake_backend import TorchBackend

pytestmark = pytest.mark.gpu
REF = TorchBackend()
DEV = "cuda"


@pytest.fixture(scope="module")
def K():
    assert torch.cuda.is_available()
    torch.backends.cuda.matmul.allow_tf32 = False
    b = ops.get()
    assert b.name == "cuda"
    return b


def _g(seed):
    return torch.Generator(device="cpu").manual_seed(seed)


def _rand(shape, dtype, seed, scale=1.0):
    return (torch.randn(shape, generator=_g(seed)) * scale).to(dtype).to(DEV)


def _check(got, ref, dtype, what="", scale=None):
    got, ref = got.float(), ref.float()
    assert got.shape == ref.shape, (what, got.shape, ref.shape)
    assert torch.isfinite(got).all(), f"{what}: non-finite output"
    s = scale if scale is not None else max(ref.abs().max().item(), 1e-6)
    err = (got - ref).abs().max().item() / s
    tol = 2e-4 if dtype == torch.float32 else 2e-2
    assert err <= tol, f"{what}: rel-to-scale err {err:.3e} > {tol}"


# ------------------------------------------------------------------------------------------- GEMM
def _gemm_case(K, M, N, Kd, tA, tB, dtype, impl, epilogue=L.EPI_NONE, bias=False, batch=0, bcast_b=False,
               out_dtype=None, seed=0):
    sa = (Kd, M) if tA else (M, Kd)
    sb = (Kd, N) if tB else (N, Kd)
    if batch:
        sa = (batch,) + sa
        if not bcast_b:
            sb = (batch,) + sb
    A = _rand(sa, dtype, seed + 1)
    B = _rand(sb, dtype, seed + 2)
    bs = _rand((N,), torch.float32, seed + 3) if bias else None
    cs = ((batch, M, N) if batch else (M, N))
    aux = None
    if epilogue == L.EPI_RESID:
        aux = _rand(cs, torch.float32, seed + 4)
        out_dtype = torch.float32
    elif epilogue in (L.EPI_QGELU_BWD, L.EPI_RELU_BWD):
        aux = _rand(cs, dtype, seed + 4)
    kw = dict(transA=tA, transB=tB, bias=bs, epilogue=epilogue, aux=aux, out_dtype=out_dtype)
    want_pre = epilogue == L.EPI_QGELU
    got = K.gemm(A, B, impl=impl, want_preact=want_pre, **kw)
    ref = REF.gemm(A, B, want_preact=want_pre, **kw)
    torch.cuda.synchronize()
    # fp32: error relative to the typical magnitude sqrt(K); bf16 outputs: one bf16 ulp (2^-8) of the largest
    # output on each side of the comparison, i.e. relative to max |ref|
    scale = math.sqrt(Kd)
    if (out_dtype or dtype) == torch.bfloat16:
        r0 = ref[0] if want_pre else ref
        scale = max(scale, r0.float().abs().max().item() / 2)
    if want_pre:
        _check(got[0], ref[0], dtype, "gemm out", scale)
        _check(got[1], ref[1], dtype, "gemm preact", scale)
    else:
        _check(got, ref, dtype, f"gemm M{M} N{N} K{Kd} tA{tA} tB{tB} epi{epilogue}", scale)


@pytest.mark.parametrize("tA,tB", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("shape", [(128, 128, 64), (256, 512, 512), (200, 72, 136), (64, 8, 40), (1, 16, 8)])
def test_gemm_simt_fp32(K, tA, tB, shape):
    _gemm_case(K, *shape, tA, tB, torch.float32, L.GEMM_SIMT, bias=True)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("epi", [L.EPI_NONE, L.EPI_QGELU, L.EPI_RESID, L.EPI_QGELU_BWD, L.EPI_RELU, L.EPI_RELU_BWD])
def test_gemm_simt_epilogues(K, dtype, epi):
    _gemm_case(K, 192, 136, 96, False, False, dtype, L.GEMM_SIMT, epilogue=epi, bias=True)


def test_gemm_simt_small_n_and_batch(K):
    _gemm_case(K, 100, 1, 512, False, False, torch.float32, L.GEMM_SIMT, bias=True)
    _gemm_case(K, 100, 2, 512, False, False, torch.bfloat16, L.GEMM_SIMT, bias=True, out_dtype=torch.float32)
    _gemm_case(K, 96, 40, 64, False, False, torch.float32, L.GEMM_SIMT, batch=3)
    _gemm_case(K, 96, 40, 64, False, False, torch.float32, L.GEMM_SIMT, batch=3, bcast_b=True)


@pytest.mark.parametrize("tA,tB", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("shape", [(128, 128, 64), (256, 256, 512), (384, 512, 2048), (8192, 1536, 512),
                                   (200, 72, 136), (130, 264, 72), (64, 8, 64)])
def test_gemm_tc_layouts(K, tA, tB, shape):
    M, N, _ = shape
    if (tA and M % 8) or (tB and N % 8):
        # an MN-major operand whose leading dimension is not a multiple of 16 bytes cannot be described to TMA:
        # the forced tensor-core path must refuse it loudly (AUTO dispatch falls back to the SIMT kernel)
        with pytest.raises(L.ExgError):
            _gemm_case(K, *shape, tA, tB, torch.bfloat16, L.GEMM_TC)
        _gemm_case(K, *shape, tA, tB, torch.bfloat16, L.GEMM_AUTO)
        return
    _gemm_case(K, *shape, tA, tB, torch.bfloat16, L.GEMM_TC)


@pytest.mark.parametrize("epi", [L.EPI_NONE, L.EPI_QGELU, L.EPI_RESID, L.EPI_QGELU_BWD, L.EPI_RELU, L.EPI_RELU_BWD])
def test_gemm_tc_epilogues(K, epi):
    _gemm_case(K, 300, 520, 512, False, False, torch.bfloat16, L.GEMM_TC, epilogue=epi, bias=True)


def test_gemm_tc_fp32_out_splitk_batch(K):
    # weight-gradient shape: tiny M x N, long K -> split-K with fp32 atomics
    _gemm_case(K, 512, 512, 8192, True, True, torch.bfloat16, L.GEMM_TC, out_dtype=torch.float32)
    _gemm_case(K, 1536, 512, 12288, True, True, torch.bfloat16, L.GEMM_TC, out_dtype=torch.float32)
    _gemm_case(K, 256, 264, 512, False, False, torch.bfloat16, L.GEMM_TC, batch=3)
    _gemm_case(K, 256, 264, 512, False, False, torch.bfloat16, L.GEMM_TC, batch=3, bcast_b=True)
    _gemm_case(K, 192, 136, 520, False, True, torch.bfloat16, L.GEMM_TC, batch=2)
    _gemm_case(K, 264, 64, 520, True, True, torch.bfloat16, L.GEMM_TC, batch=2, out_dtype=torch.float32)


def test_gemm_auto_dispatch(K):
    _gemm_case(K, 256, 512, 512, False, False, torch.bfloat16, L.GEMM_AUTO, bias=True)
    _gemm_case(K, 100, 1, 512, False, False, torch.bfloat16, L.GEMM_AUTO, bias=True, out_dtype=torch.float32)
    _gemm_case(K, 256, 512, 512, False, False, torch.float32, L.GEMM_AUTO, bias=True)


# ------------------------------------------------------------------------------------------- row-wise
@pytest.mark.parametrize("xd,yd", [(torch.float32, torch.float32), (torch.float32, torch.bfloat16),
                                   (torch.bfloat16, torch.float32)])
@pytest.mark.parametrize("C", [512, 768, 4096])
def test_layernorm_fwd_bwd(K, xd, yd, C):
    rows = 77
    x = _rand((rows, C), xd, 1, 2.0)
    g = 1 + 0.1 * _rand((C,), torch.float32, 2)
    b = 0.1 * _rand((C,), torch.float32, 3)
    add = _rand((7, C), torch.float32, 4)
    y, y2 = K.layernorm_fwd(x, g, b, out_dtype=yd, y2_dtype=torch.float32, add=add, add_period=7)
    yr, y2r = REF.layernorm_fwd(x, g, b, out_dtype=yd, y2_dtype=torch.float32, add=add, add_period=7)
    _check(y, yr, yd, "ln fwd")
    _check(y2, y2r, torch.float32, "ln fwd y2")
    dy1 = _rand((rows, C), yd, 5)
    dy2 = _rand((rows, C), yd, 6)
    dres = _rand((rows, C), torch.float32, 7)
    got = K.layernorm_bwd(dy1, x, g, dy2=dy2, dres=dres, dxT_dtype=torch.bfloat16)
    ref = REF.layernorm_bwd(dy1, x, g, dy2=dy2, dres=dres, dxT_dtype=torch.bfloat16)
    _check(got[0], ref[0], torch.float32 if xd == torch.float32 and yd == torch.float32 else torch.bfloat16, "ln dx")
    _check(got[1], ref[1], torch.bfloat16, "ln dxT")
    _check(got[2], ref[2], torch.float32 if yd == torch.float32 and xd == torch.float32 else torch.bfloat16, "ln dgamma")
    _check(got[3], ref[3], torch.float32 if yd == torch.float32 else torch.bfloat16, "ln dbeta")
    if C <= 2048:       # fused bias-gradient sink: column sums of dx accumulated into a caller buffer
        acc = _rand((C,), torch.float32, 8)
        accr = acc.clone()
        K.layernorm_bwd(dy1, x, g, dy2=dy2, dres=dres, dxT_dtype=torch.bfloat16, dxsum_out=acc)
        REF.layernorm_bwd(dy1, x, g, dy2=dy2, dres=dres, dxT_dtype=torch.bfloat16, dxsum_out=accr)
        _check(acc, accr, torch.float32 if xd == torch.float32 and yd == torch.float32 else torch.bfloat16, "ln dxsum")


@pytest.mark.parametrize("rows", [5, 8192, 12288 + 3])
def test_layernorm_bwd_pipelined_bf16(K, rows):
    """The training configuration (x fp32, dy1 / dy2 bf16, raw-bits prefetch) over many rows per warp."""
    C = 512
    x = _rand((rows, C), torch.float32, 1, 2.0)
    g = 1 + 0.1 * _rand((C,), torch.float32, 2)
    dy1, dy2 = _rand((rows, C), torch.bfloat16, 5), _rand((rows, C), torch.bfloat16, 6)
    dres = _rand((rows, C), torch.float32, 7)
    for d2 in (dy2, None):
        acc = torch.zeros(C, device=DEV)
        accr = torch.zeros(C, device=DEV)
        got = K.layernorm_bwd(dy1, x, g, dy2=d2, dres=dres, dxT_dtype=torch.bfloat16, dxsum_out=acc)
        ref = REF.layernorm_bwd(dy1, x, g, dy2=d2, dres=dres, dxT_dtype=torch.bfloat16, dxsum_out=accr)
        _check(got[0], ref[0], torch.float32, "ln dx")
        _check(got[1], ref[1], torch.bfloat16, "ln dxT")
        _check(got[2], ref[2], torch.float32, "ln dgamma")
        _check(got[3], ref[3], torch.float32, "ln dbeta")
        _check(acc, accr, torch.float32, "ln dxsum")


def test_layernorm_rowmaps(K):
    B, T, N, C = 3, 5, 2, 512
    x = _rand((B * T, C), torch.float32, 1)
    g = 1 + 0.1 * _rand((C,), torch.float32, 2)
    b = 0.1 * _rand((C,), torch.float32, 3)
    out = torch.zeros(B * (T + N), C, device=DEV)
    outr = torch.zeros(B * (T + N), C, device=DEV)
    K.layernorm_fwd(x, g, b, out_dtype=torch.float32, out=out, out_map=(T, T + N, 0))
    REF.layernorm_fwd(x, g, b, out_dtype=torch.float32, out=outr, out_map=(T, T + N, 0))
    _check(out, outr, torch.float32, "ln fwd map")
    dy = _rand((B * (T + N), C), torch.float32, 4)
    got = K.layernorm_bwd(dy, x, g, map1=(T, T + N, 0))
    ref = REF.layernorm_bwd(dy, x, g, map1=(T, T + N, 0))
    for a, r, n in zip(got, ref, ("dx", "dxT", "dg", "db")):
        if r is not None:
            _check(a, r, torch.float32, "ln bwd map " + n)


def test_add_rows_and_l2norm(K):
    B, T, N, C = 4, 6, 3, 512
    xv = _rand((B * T, C), torch.float32, 1)
    xt = _rand((B * N, C), torch.float32, 2)
    pos = _rand((T, C), torch.float32, 3)
    out = torch.zeros(B * (T + N), C, device=DEV)
    outr = torch.zeros_like(out)
    for Kb, o in ((K, out), (REF, outr)):
        Kb.add_rows(xv, B * T, add=pos, add_period=T, out=o, map_out=(T, T + N, 0))
        Kb.add_rows(xt, B * N, out=o, map_out=(N, T + N, T))
    _check(out, outr, torch.float32, "concat")
    g1, _ = K.add_rows(out, B * N, map_x=(N, T + N, T), out_dtype=torch.bfloat16)
    g2, _ = REF.add_rows(out, B * N, map_x=(N, T + N, T), out_dtype=torch.bfloat16)
    _check(g1, g2, torch.bfloat16, "gather")
    for dt in (torch.float32, torch.bfloat16):
        st = _rand((2, B * (T + N), C), dt, 5)
        y, inv = K.l2norm_fwd(st.view(-1, C), 2 * B * T, map_x=(T, T + N, 0))
        yr, invr = REF.l2norm_fwd(st.view(-1, C), 2 * B * T, map_x=(T, T + N, 0))
        _check(y, yr, dt, "l2 fwd")
        _check(inv, invr, dt, "l2 inv")
        dy = _rand(tuple(y.shape), dt, 6)
        dx, dxr = torch.zeros_like(st), torch.zeros_like(st)
        K.l2norm_bwd(dy, yr, invr, dx.view(-1, C), map_dx=(T, T + N, 0))
        REF.l2norm_bwd(dy, yr, invr, dxr.view(-1, C), map_dx=(T, T + N, 0))
        _check(dx, dxr, dt, "l2 bwd")


@pytest.mark.parametrize("shape", [(12288, 512), (777, 1536), (300, 2), (300, 1), (5, 2048), (17408, 2048), (4100, 264),
                                   (63, 8), (1000, 12)])
@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_colsum(K, shape, dt):
    x = _rand(shape, dt, 1)
    _check(K.colsum(x), REF.colsum(x), torch.float32, "colsum", scale=math.sqrt(shape[0]) * 3)
    if shape[1] >= 24 and shape[1] % 8 == 0:          # a column slice (row stride > N), as the qkv gradient views are
        v = x[:, 8:shape[1] - 8]
        _check(K.colsum(v), REF.colsum(v), torch.float32, "colsum view", scale=math.sqrt(shape[0]) * 3)


def test_cast_and_stage_mean(K):
    x = _rand((1000003,), torch.float32, 1)
    assert torch.equal(K.cast(x, torch.bfloat16), x.to(torch.bfloat16))
    assert torch.equal(K.cast(x.to(torch.bfloat16), torch.float32), x.to(torch.bfloat16).float())
    st = _rand((6, 40, 512), torch.bfloat16, 2)
    m, mT = K.stage_mean(st, outT_dtype=torch.bfloat16)
    mr, mTr = REF.stage_mean(st, outT_dtype=torch.bfloat16)
    _check(m, mr, torch.float32, "stage_mean")
    _check(mT, mTr, torch.bfloat16, "stage_mean T")


# ------------------------------------------------------------------------------------------- attention
@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,Sq,Sk,masked", [(2, 64, 64, True), (3, 96, 96, True), (2, 100, 100, True),
                                            (2, 8, 200, True), (1, 300, 300, False), (4, 64, 64, False),
                                            (2, 24, 120, True), (3, 128, 128, True), (2, 200, 72, True),
                                            (2, 1024, 1024, True), (1, 1100, 1100, True), (1, 64, 2080, True),
                                            (2, 520, 130, False), (1, 129, 257, True)])
def test_attention(K, dt, B, Sq, Sk, masked):
    H, d = 8, 64
    self_attn = Sq == Sk
    if self_attn:
        qkv = _rand((B * Sq, 3 * H * d), dt, 1)
        q, k, v = qkv[:, :512], qkv[:, 512:1024], qkv[:, 1024:]
    else:
        q = _rand((B * Sq, H * d), dt, 1)
        kv = _rand((B * Sk, 2 * H * d), dt, 2)
        k, v = kv[:, :512], kv[:, 512:]
    pad = None
    if masked:
        pad = torch.zeros(B, Sk, dtype=torch.uint8, device=DEV)
        for b in range(B):
            pad[b, Sk - 1 - 3 * b - (Sk // 3 if b == 0 else 0):] = 1
    scale = d ** -0.5
    o, lse = K.attn_fwd(q, k, v, pad, B, H, Sq, Sk, scale)
    orf, lser = REF.attn_fwd(q, k, v, pad, B, H, Sq, Sk, scale)
    _check(o, orf, dt, "attn o")
    _check(lse, lser, dt, "attn lse")
    do = _rand((B * Sq, H * d), dt, 3)
    if self_attn:
        dqkv, dqkvr = torch.zeros_like(qkv), torch.zeros_like(qkv)
        K.attn_bwd(q, k, v, orf, do, lser, pad, dqkv[:, :512], dqkv[:, 512:1024], dqkv[:, 1024:], B, H, Sq, Sk, scale)
        REF.attn_bwd(q, k, v, orf, do, lser, pad, dqkvr[:, :512], dqkvr[:, 512:1024], dqkvr[:, 1024:], B, H, Sq, Sk, scale)
        _check(dqkv, dqkvr, dt, "attn dqkv")
    else:
        dq, dkv = torch.zeros_like(q), torch.zeros_like(kv)
        dqr, dkvr = torch.zeros_like(q), torch.zeros_like(kv)
        K.attn_bwd(q, k, v, orf, do, lser, pad, dq, dkv[:, :512], dkv[:, 512:], B, H, Sq, Sk, scale)
        REF.attn_bwd(q, k, v, orf, do, lser, pad, dqr, dkvr[:, :512], dkvr[:, 512:], B, H, Sq, Sk, scale)
        _check(dq, dqr, dt, "attn dq")
        _check(dkv, dkvr, dt, "attn dkv")


def test_flash_attention_lazy_rescale_paths(K):
    """attention_flash.cu forward takes exponentials relative to the running maximum of the PREVIOUS key tiles and only
    re-bases (exact two-pass tile, O folded out of TMEM) when some row's tile maximum jumps by more than 2^20.  Keys
    whose scores jump far beyond that in later tiles, a clip whose first key tile is fully masked (running maximum
    -inf, then finite) and ordinary tiles in between: every path against the spec."""
    dt = torch.bfloat16
    B, H, d, S = 3, 8, 64, 640                      # 5 key tiles
    q = _rand((B * S, H * d), dt, 1)
    kv = _rand((B * S, 2 * H * d), dt, 2)
    k, v = kv[:, :512], kv[:, 512:]
    kb = k.view(B, S, H * d)
    kb[0, 256:384] *= 24.0                          # clip 0: tile 2 scores ~24x larger (jump), tiles 3-4 back to normal
    kb[1, 512:] *= 40.0                             # clip 1: the last tile jumps
    pad = torch.zeros(B, S, dtype=torch.uint8, device=DEV)
    pad[2, :128] = 1                                # clip 2: first key tile masked out entirely
    pad[2, 600:] = 1
    scale = d ** -0.5
    o, lse = K.attn_fwd(q, k, v, pad, B, H, S, S, scale)
    orf, lser = REF.attn_fwd(q, k, v, pad, B, H, S, S, scale)
    assert torch.isfinite(o.float()).all() and torch.isfinite(lse).all()
    _check(o, orf, dt, "flash lazy o")
    _check(lse, lser, dt, "flash lazy lse")


# ------------------------------------------------------------------------------------------- NCE
@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("S,B,T,N", [(2, 3, 16, 8), (3, 5, 24, 4), (6, 16, 64, 32), (2, 20, 5, 216), (1, 7, 3, 40),
                                     (2, 33, 7, 8), (2, 64, 64, 32)])
def test_nce(K, dt, S, B, T, N):
    Lg = (torch.rand((S, B * T, B * N), generator=_g(1)) * 2 - 1).to(dt).to(DEV)
    tgt = (torch.rand((B, N, T), generator=_g(2)) < 0.15).to(torch.uint8).to(DEV)
    cv = (torch.rand((B * N,), generator=_g(3)) < 0.8).to(torch.uint8).to(DEV)
    tgt = tgt * cv.view(B, N, 1)
    it = 1 / 0.07
    assert torch.equal(K.nce_pack_targets(tgt), REF.nce_pack_targets(tgt))
    got = K.nce_fwd(Lg, tgt, cv, B, T, N, it)
    ref = REF.nce_fwd(Lg, tgt, cv, B, T, N, it)
    for a, r, n in zip(got, ref, ("row_all", "row_pos", "col_all", "col_pos")):
        _check(a, r, torch.float32, "nce " + n)
    wr = (torch.rand((B * T,), generator=_g(4)) * (torch.rand((B * T,), generator=_g(5)) < 0.7)).to(DEV)
    wc = (torch.rand((B * N,), generator=_g(6)) * (torch.rand((B * N,), generator=_g(7)) < 0.7)).to(DEV)
    wr = wr * (ref[1][0] > 0)
    wc = wc * (ref[3][0] > 0)
    G = K.nce_bwd(Lg, tgt, cv, ref, wr, wc, B, T, N, it)
    Gr = REF.nce_bwd(Lg, tgt, cv, ref, wr, wc, B, T, N, it)
    _check(G, Gr, dt, "nce G")
    _check(K.own_max(Lg, S - 1, B, T, N), REF.own_max(Lg, S - 1, B, T, N), torch.float32, "own_max")
    # own-clip block of masked rows dropped (the self-labelling branch's in-place -6e4 fill, loss.py:100-106)
    rm = (torch.rand((B * T,), generator=_g(8)) < 0.3).to(torch.uint8).to(DEV)
    got = K.nce_fwd(Lg, tgt, cv, B, T, N, it, own_row_mask=rm)
    ref = REF.nce_fwd(Lg, tgt, cv, B, T, N, it, own_row_mask=rm)
    for a, r, n in zip(got, ref, ("row_all", "row_pos", "col_all", "col_pos")):
        _check(a, r, torch.float32, "nce masked " + n)
    wr2, wc2 = wr * (ref[1][0] > 0), wc * (ref[3][0] > 0)
    G = K.nce_bwd(Lg, tgt, cv, ref, wr2, wc2, B, T, N, it, own_row_mask=rm)
    _check(G, REF.nce_bwd(Lg, tgt, cv, ref, wr2, wc2, B, T, N, it, own_row_mask=rm), dt, "nce masked G")
    # weighted rows / columns whose every positive was dropped: the dropped positives carry -w/#entries
    G = K.nce_bwd(Lg, tgt, cv, ref, wr, wc, B, T, N, it, own_row_mask=rm)
    Gr = REF.nce_bwd(Lg, tgt, cv, ref, wr, wc, B, T, N, it, own_row_mask=rm)
    assert (wr * (ref[1][0] <= 0)).any(), "case does not exercise the leak term"
    _check(G, Gr, dt, "nce masked leak G")


@pytest.mark.parametrize("th,head,use_abs,nce_w", [(0.5, 1, 0, 1), (0.3, 0, 0, 1), (0.0, 1, 1, 1), (0.5, 1, 1, 0),
                                                   (0.7, 1, 0, 1)])
@pytest.mark.parametrize("E,D,B,T,N", [(2, 3, 3, 16, 8), (6, 6, 128, 64, 32), (1, 2, 37, 9, 20), (3, 1, 512, 4, 40)])
def test_tan_loss_tail(K, E, D, B, T, N, th, head, use_abs, nce_w):
    """exg_nce_logdiff + exg_tan_loss_tail against the reference-style torch spec (boolean compaction,
    torch.quantile, Tensor.std, F.binary_cross_entropy_with_logits): every logged scalar, the row / column weights
    of the backward and d loss / d alignability logits."""
    it = 1 / 0.07
    tgt = (torch.rand((B, N, T), generator=_g(2)) < 0.15).to(torch.uint8).to(DEV)
    cv = (torch.rand((B * N,), generator=_g(3)) < 0.8).to(torch.uint8).to(DEV)
    tgt = (tgt * cv.view(B, N, 1)).contiguous()
    sums = []
    for S, seed in ((E, 11), (D, 12)):
        Lg = (torch.rand((S, B * T, B * N), generator=_g(seed)) * 2 - 1).to(DEV)
        sums.append((REF.nce_fwd(Lg, tgt, cv, B, T, N, it), REF.own_max(Lg, S - 1, B, T, N).contiguous()))
        del Lg
    def pack(sm):      # the [2,S,R] / [2,S,Cc] storage exg_nce_fwd writes
        rows, cols = torch.stack((sm[0], sm[1])).contiguous(), torch.stack((sm[2], sm[3])).contiguous()
        return rows[0], rows[1], cols[0], cols[1]
    (sd, md), (sj, mj) = sums
    sd, sj = pack(sd), pack(sj)
    la = (torch.randn((B * N,), generator=_g(13)) * 2).to(DEV) if head else None
    abs_c = torch.rand((B * N,), generator=_g(14)).to(DEV) if (use_abs and head) else None
    out, w = K.tan_loss_tail(tgt, cv, sd, sj, md, mj, la, abs_c, B, T, N, it, th, head, nce_w)
    ro, rw = REF.tan_loss_tail(tgt, cv, sd, sj, md, mj, la, abs_c, B, T, N, it, th, head, nce_w)
    names = ("loss", "loss-dual", "loss-joint", "loss-dual-all", "loss-joint-all", "loss-total", "bce", "top1",
             "n_rows", "n_cols", "n_valid")
    for i, n in enumerate(names):
        a, r = float(out[i]), float(ro[i])
        assert abs(a - r) <= 2e-5 * max(1.0, abs(r)), (n, a, r)
    for i, n in enumerate(("wr_d", "wc_d", "wr_j", "wc_j", "dla")):
        assert torch.equal(w[i] != 0, rw[i] != 0) or n == "dla", n       # identical keep / row masks
        _check(w[i], rw[i], torch.float32, "tail " + n)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("S,W,B,T,N", [(2, 2, 3, 16, 8), (3, 4, 16, 64, 32), (2, 3, 5, 7, 12), (1, 8, 16, 64, 32)])
def test_nce_row_blocks(K, dt, S, W, B, T, N):
    """Cross-GPU negatives: every rank holds the row block [S, B*T, W*B*N] of the global logits (Bc, clip0).  Row
    sums are the global ones' slices, column sums add up over ranks, and with the summed column sums the backward
    of a block is the slice of the global gradient."""
    Bg = W * B
    Lg = (torch.rand((S, Bg * T, Bg * N), generator=_g(1)) * 2 - 1).to(dt).to(DEV)
    tgt = (torch.rand((Bg, N, T), generator=_g(2)) < 0.15).to(torch.uint8).to(DEV)
    cv = (torch.rand((Bg * N,), generator=_g(3)) < 0.8).to(torch.uint8).to(DEV)
    tgt = tgt * cv.view(Bg, N, 1)
    rm = (torch.rand((Bg * T,), generator=_g(8)) < 0.2).to(torch.uint8).to(DEV)
    it = 1 / 0.07
    for mask in (None, rm):
        ref = REF.nce_fwd(Lg, tgt, cv, Bg, T, N, it, own_row_mask=mask)
        wr = (torch.rand((Bg * T,), generator=_g(4)) * (torch.rand((Bg * T,), generator=_g(5)) < 0.7)).to(DEV)
        wc = (torch.rand((Bg * N,), generator=_g(6)) * (torch.rand((Bg * N,), generator=_g(7)) < 0.7)).to(DEV)
        Gr = REF.nce_bwd(Lg, tgt, cv, ref, wr, wc, Bg, T, N, it, own_row_mask=mask)
        mref = REF.own_max(Lg, S - 1, Bg, T, N)
        ca, cp, parts = 0, 0, []
        for r in range(W):
            rows = slice(r * B * T, (r + 1) * B * T)
            blk = Lg[:, rows].contiguous()
            m_loc = mask[rows].contiguous() if mask is not None else None
            got = K.nce_fwd(blk, tgt[r * B:(r + 1) * B].contiguous(), cv, B, T, N, it, own_row_mask=m_loc, Bc=Bg,
                            clip0=r * B)
            _check(got[0], ref[0][:, rows], torch.float32, "block row_all")
            _check(got[1], ref[1][:, rows], torch.float32, "block row_pos")
            ca, cp = ca + got[2], cp + got[3]
            parts.append((blk, got, m_loc))
            _check(K.own_max(blk, S - 1, B, T, N, Bc=Bg, clip0=r * B), mref[r * B:(r + 1) * B], torch.float32,
                   "block own_max")
        _check(ca, ref[2], torch.float32, "summed col_all")
        _check(cp, ref[3], torch.float32, "summed col_pos")
        for r, (blk, got, m_loc) in enumerate(parts):
            rows = slice(r * B * T, (r + 1) * B * T)
            G = K.nce_bwd(blk, tgt[r * B:(r + 1) * B].contiguous(), cv, (ref[0][:, rows].contiguous(),
                          ref[1][:, rows].contiguous(), ref[2], ref[3]), wr[rows].contiguous(), wc, B, T, N, it,
                          own_row_mask=m_loc, Bc=Bg, clip0=r * B)
            _check(G, Gr[:, rows], dt, "block G")


# ------------------------------------------------------------------------------------------- optimizer
def test_adamw_and_ema(K):
    n = 100003
    p, g = _rand((n + 1,), torch.float32, 1)[:n], _rand((n + 1,), torch.float32, 2)[:n]
    m, v = torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    sh = torch.zeros(n, dtype=torch.bfloat16, device=DEV)
    pr, mr, vr, shr = p.clone(), m.clone(), v.clone(), sh.clone()
    for step in (1, 2, 3):
        K.adamw(p, g, m, v, sh, 1e-3, 0.9, 0.999, 1e-8, 1e-2, step, 0.5)
        REF.adamw(pr, g, mr, vr, shr, 1e-3, 0.9, 0.999, 1e-8, 1e-2, step, 0.5)
    _check(p, pr, torch.float32, "adamw p")
    _check(m, mr, torch.float32, "adamw m")
    _check(v, vr, torch.float32, "adamw v")
    _check(sh, shr, torch.bfloat16, "adamw shadow")
    opt_p = torch.nn.Parameter(_rand((n + 1,), torch.float32, 1)[:n].clone())
    opt = torch.optim.AdamW([opt_p], lr=1e-3, weight_decay=1e-2)
    for _ in range(3):
        opt_p.grad = g * 0.5
        opt.step()
    _check(p, opt_p.detach(), torch.float32, "adamw vs torch.optim.AdamW")
    t, o = _rand((n,), torch.float32, 3), _rand((n,), torch.float32, 4)
    tr = t.clone()
    K.ema(t, o, 0.999)
    REF.ema(tr, o, 0.999)
    _check(t, tr, torch.float32, "ema")


@pytest.mark.parametrize("M,Kd,bias", [(8192, 512, True), (12288, 2048, True), (200, 512, False), (129, 136, True),
                                       (20000, 512, True), (64, 64, True)])
def test_gemm_add_ln(K, M, Kd, bias):
    """exg_gemm_add_ln (out-projection + residual add + LayerNorm in one tcgen05 kernel) vs GEMM -> add -> LayerNorm in
    fp32 torch: the fp32 residual stream to 2e-3 of its scale (bf16 operands, fp32 accumulate), the bf16 normalised
    rows to bf16 resolution; row tails (M % 128), K tails (K % 64), more tiles than SMs."""
    A = _rand((M, Kd), torch.bfloat16, 1)
    W = _rand((512, Kd), torch.bfloat16, 2) * (Kd ** -0.5)
    b = _rand((512,), torch.float32, 3) if bias else None
    resid = _rand((M, 512), torch.float32, 4) * 3 + 0.5
    g = 1 + 0.1 * _rand((512,), torch.float32, 5)
    be = 0.1 * _rand((512,), torch.float32, 6)
    y, s_ = K.gemm_add_ln(A, W, b, resid, g, be)          # (the engine takes this path with EXG_GEMM_LN=1)
    yr, sr = REF.gemm_add_ln(A, W, b, resid, g, be)
    assert y.dtype == torch.bfloat16 and s_.dtype == torch.float32
    err = (s_ - sr).abs().max().item()
    assert err <= 2e-3 * sr.abs().max().item(), f"s: {err}"
    # y from the kernel's own s (isolates the LayerNorm part from GEMM rounding)
    y_own = torch.nn.functional.layer_norm(s_, (512,), g, be, 1e-5)
    assert (y.float() - y_own).abs().max().item() <= 2e-2 * y_own.abs().max().item()
    assert (y.float() - yr.float()).abs().max().item() <= 3e-2 * yr.float().abs().max().item()
    out = torch.empty_like(y)
    y2, _ = K.gemm_add_ln(A, W, b, resid, g, be, out=out)
    assert y2.data_ptr() == out.data_ptr() and torch.equal(out, y)


def test_align_finish(K):
    """exg_align_finish vs the reference's stitch loop (eval_zeroshot_align.py:200-222): overlapping windows, sentences
    covered by 1-3 windows, a sentence covered by none (all-zero row -> -6e4 everywhere), fp32 exact argmax."""
    from exoground_b200.eval.align import window_plan
    g = torch.Generator().manual_seed(11)
    vlen, Kn, seq = 301, 37, 64
    start = sorted((torch.rand(Kn, generator=g) * (vlen - 10)).tolist())
    end = [s_ + 2 + 5 * float(torch.rand(1, generator=g)) for s_ in start]
    aligned = [0 if i % 3 == 0 else 1 for i in range(Kn)]      # windows are anchored on the non-aligned sentences
    plan = window_plan(vlen, seq, start, end, aligned)
    assert len(plan) > 4
    T = max(p_[1] for p_ in plan)
    N = max(p_[3] - p_[2] + 1 for p_ in plan)
    sj = torch.randn(len(plan), T, N, generator=g).to(DEV) * 3
    sd = torch.randn(len(plan), T, N, generator=g).to(DEV) * 3
    win = torch.tensor(plan, dtype=torch.int32, device=DEV)
    got = K.align_finish(sj, sd, win, Kn, vlen)
    ref = REF.align_finish(sj, sd, win, Kn, vlen)
    _check(got[0], ref[0], torch.float32, "align sim")
    assert torch.allclose(got[1], ref[1], rtol=1e-4, atol=1e-7), "align prob"
    assert torch.equal(got[2].cpu(), ref[2].cpu()), "align pred"
    _check(got[3], ref[3], torch.float32, "align max")
    covered = torch.zeros(Kn, dtype=torch.bool)
    for s_, ln, l, r in plan:
        covered[l:r + 1] = True
    if (~covered).any():
        assert float(got[0][~covered].max()) == -6e4


def test_per_parameter_clip(K):
    """exg_grad_sumsq / exg_grad_clip_scale / exg_adamw_clip vs utils/train_utils.py:3-13 + torch.optim.AdamW."""
    sizes = [512, 1536 * 512, 70, 2048 * 512 + 3, 64, 4096 * 512, 1]
    offs, off = [], 0
    for n in sizes:
        offs.append(off)
        off += (n + 63) // 64 * 64
    total = off
    tab = torch.zeros(total // 64, dtype=torch.int32)
    for i, (o, n) in enumerate(zip(offs, sizes)):
        tab[o // 64:(o + n + 63) // 64] = i
    tab = tab.to(DEV)
    g = torch.zeros(total, device=DEV)
    gen = torch.Generator(device="cpu").manual_seed(5)
    for i, (o, n) in enumerate(zip(offs, sizes)):
        g[o:o + n] = (torch.randn(n, generator=gen) * (10.0 ** (i % 3 - 2))).to(DEV)
    clip, gscale = 0.7, 0.5
    ss = K.grad_sumsq(g, tab, len(sizes))
    want = torch.stack([g[o:o + n].double().pow(2).sum() for o, n in zip(offs, sizes)]).float()
    _check(ss, want, torch.float32, "grad_sumsq")
    _check(ss, REF.grad_sumsq(g, tab, len(sizes)), torch.float32, "grad_sumsq vs spec")
    # reference semantics: clip the (mean) gradient in place, then AdamW
    ps = [torch.nn.Parameter(torch.randn(n, generator=gen).to(DEV)) for n in sizes]
    p_flat = torch.zeros(total, device=DEV)
    for o, n, p in zip(offs, sizes, ps):
        p_flat[o:o + n] = p.detach()
    opt = torch.optim.AdamW(ps, lr=1e-3, weight_decay=1e-2)
    m, v = torch.zeros(total, device=DEV), torch.zeros(total, device=DEV)
    sh = torch.zeros(total, dtype=torch.bfloat16, device=DEV)
    g2 = g.clone()
    for step in (1, 2):
        for o, n, p in zip(offs, sizes, ps):
            p.grad = g[o:o + n] * gscale
            c = clip / (p.grad.norm(2) + 1e-6)
            if c < 1:
                p.grad.mul_(c)
        opt.step()
        K.grad_sumsq(g, tab, len(sizes), out=ss)
        K.adamw_clip(p_flat, g, m, v, sh, 1e-3, 0.9, 0.999, 1e-8, 1e-2, step, tab, ss, clip, grad_scale=gscale)
    for o, n, p in zip(offs, sizes, ps):
        _check(p_flat[o:o + n], p.detach(), torch.float32, f"adamw_clip segment {n}")
    _check(sh, p_flat.bfloat16(), torch.bfloat16, "adamw_clip shadow")
    assert torch.equal(g, g2)                          # the fused form leaves .grad unscaled
    K.grad_clip_scale(g2, tab, ss, clip, gscale)       # the in-place form
    for o, n in zip(offs, sizes):
        c = min(1.0, clip / (float(g[o:o + n].norm()) * gscale + 1e-6))
        _check(g2[o:o + n], g[o:o + n] * c, torch.float32, f"clip_scale segment {n}")
    # a segment view that starts past the first parameter (the decay segment of a flat buffer)
    lo = offs[3]
    p3, m3, v3 = p_flat[lo:].clone(), torch.zeros(total - lo, device=DEV), torch.zeros(total - lo, device=DEV)
    pr3, mr3, vr3 = p3.clone(), m3.clone(), v3.clone()
    K.adamw_clip(p3, g[lo:], m3, v3, None, 1e-3, 0.9, 0.999, 1e-8, 0.0, 1, tab[lo // 64:], ss, clip, grad_scale=gscale)
    REF.adamw_clip(pr3, g[lo:], mr3, vr3, None, 1e-3, 0.9, 0.999, 1e-8, 0.0, 1, tab[lo // 64:], ss, clip, grad_scale=gscale)
    _check(p3, pr3, torch.float32, "adamw_clip offset segment")


# ------------------------------------------------------------------------------------------- persistent GEMM
@pytest.mark.parametrize("tA,tB", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("shape", [(128, 128, 64), (256, 256, 512), (384, 512, 2048), (8192, 1536, 512),
                                   (200, 72, 136), (136, 264, 72), (20000, 512, 512)])
def test_gemm_tc2_layouts(K, tA, tB, shape):
    _gemm_case(K, *shape, tA, tB, torch.bfloat16, L.GEMM_TC2)


@pytest.mark.parametrize("epi", [L.EPI_NONE, L.EPI_QGELU, L.EPI_RELU, L.EPI_QGELU_BWD, L.EPI_RELU_BWD])
@pytest.mark.parametrize("shape", [(300, 520, 512), (12288, 2048, 512), (8192, 2048, 512)])
def test_gemm_tc2_epilogues(K, epi, shape):
    _gemm_case(K, *shape, False, False, torch.bfloat16, L.GEMM_TC2, epilogue=epi, bias=True)


def test_gemm_tc2_fp32_out_splitk_batch(K):
    _gemm_case(K, 512, 512, 8192, True, True, torch.bfloat16, L.GEMM_TC2, out_dtype=torch.float32)
    _gemm_case(K, 1536, 512, 12288, True, True, torch.bfloat16, L.GEMM_TC2, out_dtype=torch.float32)
    _gemm_case(K, 300, 520, 512, False, False, torch.bfloat16, L.GEMM_TC2, out_dtype=torch.float32, bias=True)
    _gemm_case(K, 256, 264, 512, False, False, torch.bfloat16, L.GEMM_TC2, batch=3)
    _gemm_case(K, 256, 264, 512, False, False, torch.bfloat16, L.GEMM_TC2, batch=3, bcast_b=True)
    _gemm_case(K, 192, 136, 520, False, True, torch.bfloat16, L.GEMM_TC2, batch=2)
    _gemm_case(K, 264, 64, 520, True, True, torch.bfloat16, L.GEMM_TC2, batch=2, out_dtype=torch.float32)
    _gemm_case(K, 512, 512, 8192, True, True, torch.bfloat16, L.GEMM_TC2, batch=2, out_dtype=torch.float32)


def test_gemm_tc2_refuses_aux_epilogues(K):
    with pytest.raises(L.ExgError):
        _gemm_case(K, 256, 512, 512, False, False, torch.bfloat16, L.GEMM_TC2, epilogue=L.EPI_RESID)


# ------------------------------------------------------------------------------------------- ExoGround ops
@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_bcast_rows(K, dt):
    g = _rand((37, 512), torch.float32, 5)
    got = K.bcast_rows(g, 3, 1.0 / 3, dt)
    _check(got, REF.bcast_rows(g, 3, 1.0 / 3, dt), dt, "bcast_rows")


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,T", [(3, 16), (5, 64), (2, 100)])
def test_xent_diag(K, dt, B, T):
    lg = _rand((B * T, T), dt, 11, scale=3.0)
    loss, lse = K.xent_diag_fwd(lg, T)
    rl, rlse = REF.xent_diag_fwd(lg, T)
    _check(loss, rl, torch.float32, "xent loss")
    _check(lse, rlse, torch.float32, "xent lse")
    G = K.xent_diag_bwd(lg, lse, 0.37, T, g_dtype=dt)
    _check(G, REF.xent_diag_bwd(lg, rlse, 0.37, T, g_dtype=dt), dt, "xent G", scale=0.37)


@pytest.mark.parametrize("same,only", [(False, False), (True, False), (True, True)])
@pytest.mark.parametrize("D", [48, 512, 4096])
def test_xview_nce(K, same, only, D):
    B, V, T = 3, 5, 16
    o = _rand((B, T, D), torch.float32, 1)
    ego = _rand((B, V, T, D), torch.float32, 2)
    g = _g(3)
    pi = torch.randint(0, V, (B, T), generator=g).to(DEV)
    ni = torch.randint(0, V, (B, T), generator=g).to(DEV)
    sf = _rand((B, T, D), torch.float32, 4) if same else None
    si = torch.randint(0, T, (B, T), generator=g).to(DEV) if same else None
    got = K.xview_nce_fwd(o, ego, pi, ni, sf, si, only, 10.0)
    ref = REF.xview_nce_fwd(o, ego, pi, ni, sf, si, only, 10.0)
    for a, b, w in zip(got[:4], ref[:4], ("loss", "l1", "pos_cos", "neg_cos")):
        _check(a, b, torch.float32, f"xview {w}")
    gl = _rand((B, T), torch.float32, 6)
    go = K.xview_nce_bwd(o, ego, pi, ni, sf, si, only, 10.0, got[4], gl)
    _check(go, REF.xview_nce_bwd(o, ego, pi, ni, sf, si, only, 10.0, None, gl), torch.float32, "xview grad")


@pytest.mark.parametrize("use_cd", [True, False])
@pytest.mark.parametrize("M", [24, 1000, 5000])
def test_ground_loss(K, use_cd, M):
    g = _g(9)
    ga = torch.rand(M, generator=g).to(DEV)
    gb = (torch.rand(M, generator=g) * 0.3).to(DEV)
    if not use_cd:
        ga, gb = ga - gb / 2, ga + gb / 2
    preds = torch.stack((ga, gb), 1) + _rand((M, 2), torch.float32, 10, scale=0.05)
    pad = (torch.rand(M, generator=g) < 0.3).to(torch.uint8).to(DEV)
    th = (0.1, 0.3, 0.5, 0.7)
    stats, iou, dp = K.ground_loss(preds.contiguous(), ga, gb, pad, use_cd, 1e-8, th)
    rs, ri, rd = REF.ground_loss(preds, ga, gb, pad, use_cd, 1e-8, th)
    _check(stats[:3], rs[:3], torch.float32, "ground stats")
    assert torch.equal(stats[3:], rs[3:]), (stats, rs)
    _check(iou, ri, torch.float32, "iou")
    _check(dp, rd, torch.float32, "dpreds")


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_rowdot_outer_wcolsum(K, dt):
    rows, C = 1037, 512
    x = _rand((rows, C), dt, 1)
    w = _rand((C,), torch.float32, 2)
    b = _rand((1,), torch.float32, 3)
    _check(K.rowdot(x, w, b), REF.rowdot(x, w, b), torch.float32, "rowdot", scale=math.sqrt(C))
    _check(K.rowdot(x, w), REF.rowdot(x, w), torch.float32, "rowdot nobias", scale=math.sqrt(C))
    g = _rand((rows,), torch.float32, 4)
    _check(K.outer_rows(g, w, dt), REF.outer_rows(g, w, dt), dt, "outer_rows")
    _check(K.wcolsum(x, g), REF.wcolsum(x, g), torch.float32, "wcolsum", scale=math.sqrt(rows))
    acc = _rand((C,), torch.float32, 5)
    acc2 = acc.clone()
    _check(K.wcolsum(x, g, out=acc), REF.wcolsum(x, g, out=acc2), torch.float32, "wcolsum acc", scale=math.sqrt(rows))
