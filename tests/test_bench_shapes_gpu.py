"""Parity at the BENCHMARK's own sizes (BASELINE.json configs[1]: Seva 1.3 B, T=21 @576x576, CFG batch 42) so that
the shapes that carry bench.py's number are checked by the driver's ``pytest -m gpu`` run, not only by a tool:

  * the full model, one teacher-forced Euler-EDM step at T=21 @72x72 latents against the fp32 oracle executed on
    the same GPU with TF32 off (step-0 denoiser output <= 2e-2, north_star), plus the RAW network output at sigma
    indices 999 / 499 / 19 (the denoiser output leans on c_out -> 0 at small sigma; the raw output does not);
  * every hot kernel at its largest in-step shape against a plain PyTorch fp32 statement of the same op:
    attention (2,10,27216) with its 80-key tail and (42,5,5184), the weight-stationary GEMMs 217728 x {320,960} x 320,
    GEGLU 217728 x 2560 x 320, conv3x3 42 x 72^2 x 320 -> 320, GroupNorm(+SiLU+Plucker modulation) 42 x 72^2 x 960,
    LayerNorm 217728 x 320 and the time-mix attention at (2, 21, 5184, 5).

Tolerances: rel-L2 <= 5e-3 for tensor-core ops (bf16 operands, fp32 accumulation, bf16 output rounding 2^-9),
<= 4e-3 for norm kernels, <= 2e-2 for the full-model step (BASELINE.json north_star).
"""
import math

import pytest
import torch
import torch.nn.functional as F

from oracle import seva_oracle as O
from oracle import synth

pytestmark = pytest.mark.gpu

T21, HW72 = 21, 72
N42 = 2 * T21
M_DS1 = N42 * HW72 * HW72  # 217728 token rows at ds1


def rel_l2(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


def bf(x):
    return x.to(torch.bfloat16)


@pytest.fixture(scope="module")
def ops():
    from seva_b200 import ops as o

    o.check_device()
    return o


@pytest.fixture(autouse=True)
def _no_tf32():
    a, b = torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    yield
    torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = a, b
    torch.cuda.empty_cache()


# ---------------------------------------------------------------------------------------------------
# full model at the benchmark configuration
# ---------------------------------------------------------------------------------------------------
def test_config2_T21_full_model_step_vs_fp32_oracle_on_gpu():
    """BASELINE configs[1] shape: 1.3 B model, T=21 (1 input + 20 targets) @72x72 latents, MultiviewCFG(1.2), scale 2.
    One Euler-EDM step through the public sampler; the CFG batch our denoiser saw is replayed through the fp32 oracle
    on the same GPU (teacher-forced), so only kernel error is measured."""
    from seva_b200 import geometry as G
    from seva_b200 import sampling as S
    from seva_b200.model import Seva, SevaParams, SGMWrapper

    dev = torch.device("cuda", 0)
    T, h, w = T21, HW72, HW72
    P = SevaParams()
    with torch.device("meta"):
        m = Seva(P)
    sd = synth.fill_state_dict({k: tuple(v.shape) for k, v in m.state_dict().items()}, seed=0)
    m.load_state_dict(sd, assign=True)
    model = SGMWrapper(m.eval()).to(dev)
    sd = {k: v.to(dev) for k, v in sd.items()}
    plk = lambda s, e, K, hh, ww: G.get_plucker_coordinates(s, e, K, target_size=(hh, ww))
    inp = synth.make_inputs(T, 1, h, w, plk, seed=1, num_steps=1)
    cu = lambda d: {k: v.to(dev) for k, v in d.items()}

    disc = S.DDPMDiscretization()
    denoiser = S.DiscreteDenoiser(disc, num_idx=1000, device=dev)
    sampler = S.EulerEDMSampler(disc, S.MultiviewCFG(1.2), num_steps=50, s_churn=0.0, s_tmin=0.0, s_tmax=999.0,
                                s_noise=1.0, verbose=False, device=dev)
    seen = {}

    def den_fn(x_in, sigma, c):
        seen["x"], seen["sigma"] = x_in.clone(), sigma.clone()
        seen["out"] = denoiser(model, x_in, sigma, c, num_frames=T)
        return seen["out"]

    sig = disc(50, device=dev)
    ones = torch.ones(T, device=dev)
    c, uc = cu(inp["c"]), cu(inp["uc"])
    eps0 = inp["eps"][0].to(dev)
    orig = torch.randn_like
    torch.randn_like = lambda t, *a, **k: eps0.to(t)
    try:
        with torch.inference_mode(), torch.autocast("cuda"):
            x0 = inp["x0"].to(dev) * torch.sqrt(1.0 + sig[0] ** 2)
            x1 = sampler.sampler_step(ones * sig[0], ones * sig[1], den_fn, x0, 2.0, c, uc, 0.0, c2w=inp["c2w"].to(dev),
                                      K=inp["K"].to(dev), input_frame_mask=inp["mask"].to(dev))
    finally:
        torch.randn_like = orig
    torch.cuda.synchronize()
    assert x1.shape == (T, 4, h, w) and torch.isfinite(x1).all()

    table = O.denoiser_table().to(dev)
    net = lambda xx, idx, cc: O.sgm_forward(sd, P, xx, idx, cc, num_frames=T)
    cc = {k: torch.cat((uc[k], c[k]), 0) for k in c}
    with torch.inference_mode():
        d_ref = O.denoise(net, table, seen["x"], seen["sigma"], dict(cc))
        e_den = rel_l2(seen["out"], d_ref)
        # the Euler update of the same step, from the oracle's denoiser output
        fs = O.multiview_scale(2.0, 1.2, inp["c2w"], inp["K"], inp["mask"]).to(dev)
        x1_ref, _ = O.sampler_step(net, table, inp["x0"].to(dev) * torch.sqrt(1.0 + sig[0] ** 2), ones * sig[0],
                                   ones * sig[1], fs, c, uc, eps0)
        e_x1 = rel_l2(x1, x1_ref)
        # raw network output (no c_out / c_skip) at three noise levels, input scaled as the denoiser would
        raw = {}
        net_cond = {k: cc[k] for k in ("crossattn", "concat", "dense_vector")}
        g = torch.Generator(device=dev).manual_seed(11)
        for idx in (999, 499, 19):
            xin = torch.randn(2 * T, 4, h, w, device=dev, generator=g)  # input * c_in has unit variance at every sigma
            t = torch.full((2 * T,), idx, device=dev, dtype=torch.int64)
            with torch.autocast("cuda"):
                ours = model(xin, t, dict(net_cond), num_frames=T)
            ref = O.sgm_forward(sd, P, xin, t, dict(net_cond), num_frames=T)
            raw[idx] = rel_l2(ours, ref)
    print(f"config2 T=21 @72x72: step-0 denoiser rel-L2 {e_den:.3e}, x1 rel-L2 {e_x1:.3e}, raw network rel-L2 "
          + ", ".join(f"idx{k}: {v:.3e}" for k, v in raw.items()))
    assert e_den < 2e-2, e_den
    assert e_x1 < 2e-2, e_x1
    for k, v in raw.items():
        assert v < 2e-2, (k, v)
    model.module.invalidate()


# ---------------------------------------------------------------------------------------------------
# attention at the two dominant in-step shapes (62 % of the attention time)
# ---------------------------------------------------------------------------------------------------
def _attn_rows_ref(qkv, B, L, heads, rows):
    """fp32 softmax(q k^T / 8) v for the query rows ``rows`` of every (batch, head): [B, len(rows), heads*64]."""
    Cc = heads * 64
    x = qkv.reshape(B, L, 3, heads, 64)
    out = torch.empty(B, len(rows), heads, 64, device=qkv.device, dtype=torch.float32)
    for b in range(B):
        for hh in range(heads):
            q = x[b, rows, 0, hh].float()
            k = x[b, :, 1, hh].float()
            v = x[b, :, 2, hh].float()
            out[b, :, hh] = torch.softmax(q @ k.t() * 0.125, dim=-1) @ v
    return out.reshape(B, len(rows), Cc)


@pytest.mark.parametrize("B,heads,L", [(2, 10, 27216), (42, 5, 5184)])
def test_attention_bench_shapes(ops, B, heads, L):
    """attn3d ds2 (27216 = 212*128 + 80: ragged key tail AND ragged last query tile) and attn2d ds1.  All keys, a
    subset of 640 query rows per (batch, head) that includes the first / last tiles (fp32 reference by rows)."""
    g = torch.Generator(device="cuda").manual_seed(L)
    Cc = heads * 64
    qkv = bf(torch.randn(B * L, 3 * Cc, device="cuda", generator=g))
    out = ops.attention_qkv(qkv, B, L, heads).reshape(B, L, Cc)
    torch.cuda.synchronize()
    rows = torch.cat([torch.arange(0, 128), torch.arange(L - 200, L),
                      torch.randint(128, L - 200, (312,), generator=torch.Generator().manual_seed(1))]).cuda()
    nb = B if B <= 2 else 3
    bsel = [0, B - 1] if B <= 2 else [0, B // 2, B - 1]
    ref = _attn_rows_ref(qkv.reshape(B, L, 3 * Cc)[bsel].reshape(nb * L, 3 * Cc), nb, L, heads, rows)
    got = out[bsel][:, rows]
    assert torch.isfinite(out).all()
    assert rel_l2(got, ref) < 5e-3, rel_l2(got, ref)


def test_timemix_attention_bench_shape(ops):
    B, T, S, heads = 2, T21, HW72 * HW72, 5
    g = torch.Generator(device="cuda").manual_seed(5)
    Cc = heads * 64
    qkv = bf(torch.randn(B * T * S, 3 * Cc, device="cuda", generator=g))
    out = ops.timemix_attention(qkv, B, T, S, heads)
    x = qkv.reshape(B, T, S, 3, heads, 64).permute(3, 0, 2, 4, 1, 5).float()  # [3, B, S, heads, T, 64]
    ref = F.scaled_dot_product_attention(x[0], x[1], x[2])                      # [B, S, heads, T, 64]
    ref = ref.permute(0, 3, 1, 2, 4).reshape(B * T * S, Cc)
    assert rel_l2(out, ref) < 4e-3, rel_l2(out, ref)


# ---------------------------------------------------------------------------------------------------
# GEMM / conv at ds1 (M = 217728 token rows, 229992 padded rows)
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("N,residual", [(320, True), (320, False), (960, False)])
def test_gemm_ds1_k320(ops, N, residual):
    """The K = 320 linears of the ds1 transformer sites (weight-stationary walk at full size): to_out / proj with
    bias + per-image bias + residual, and the fused q|k|v projection."""
    g = torch.Generator(device="cuda").manual_seed(N + residual)
    M, K, rpi = M_DS1, 320, HW72 * HW72
    A = bf(torch.randn(M, K, device="cuda", generator=g))
    W = bf(torch.randn(N, K, device="cuda", generator=g) / math.sqrt(K))
    b = torch.randn(N, device="cuda", generator=g)
    res = bf(torch.randn(M, N, device="cuda", generator=g)) if residual else None
    rb = torch.randn(M // rpi, N, device="cuda", generator=g) if residual else None
    out = ops.linear(A, W, b, residual=res, rowbias=rb, rowbias_rpi=rpi if residual else 0)
    ref = A.float() @ W.float().t() + b
    if residual:
        ref += res.float() + rb.repeat_interleave(rpi, 0)
    assert rel_l2(out, ref) < 5e-3, rel_l2(out, ref)


def test_gemm_ds1_ff(ops):
    """FeedForward at ds1: GEGLU 217728 x 2560 x 320 then the down projection 217728 x 320 x 1280 + residual."""
    from seva_b200 import pack

    g = torch.Generator(device="cuda").manual_seed(21)
    M, K, inner = M_DS1, 320, 1280
    A = bf(torch.randn(M, K, device="cuda", generator=g))
    W1 = bf(torch.randn(2 * inner, K, device="cuda", generator=g) / math.sqrt(K))
    b1 = torch.randn(2 * inner, device="cuda", generator=g) * 0.1
    bn = ops.pick_bn(2 * inner, True)
    W1p, b1p = pack.pack_geglu(W1, b1, bn)
    hid = ops.linear(A, W1p, b1p, geglu=True, bn=bn)
    y = A.float() @ W1.float().t() + b1
    x, gate = y.chunk(2, dim=-1)
    ref = x * F.gelu(gate)
    del y, x, gate
    assert hid.shape == (M, inner)
    assert rel_l2(hid, ref) < 5e-3, rel_l2(hid, ref)
    del ref
    W2 = bf(torch.randn(K, inner, device="cuda", generator=g) / math.sqrt(inner))
    b2 = torch.randn(K, device="cuda", generator=g)
    out = ops.linear(hid, W2, b2, residual=A)
    ref2 = hid.float() @ W2.float().t() + b2 + A.float()
    assert rel_l2(out, ref2) < 5e-3, rel_l2(out, ref2)


def test_conv3x3_ds1_bench_shape(ops):
    """ResBlock conv at ds1: 42 x 72^2 x 320 -> 320 with bias, time-embedding row bias and skip residual."""
    from seva_b200 import pack

    g = torch.Generator(device="cuda").manual_seed(9)
    n_img, H, Wd, Cin, Cout = N42, HW72, HW72, 320, 320
    x = bf(torch.randn(n_img, Cin, H, Wd, device="cuda", generator=g))
    w = torch.randn(Cout, Cin, 3, 3, device="cuda", generator=g) / math.sqrt(9 * Cin)
    b = torch.randn(Cout, device="cuda", generator=g)
    rb = torch.randn(n_img, Cout, device="cuda", generator=g)
    xp = F.pad(x.permute(0, 2, 3, 1), (0, 0, 1, 1, 1, 1)).reshape(-1, Cin).contiguous()
    res = F.pad(bf(torch.randn(n_img, H, Wd, Cout, device="cuda", generator=g)), (0, 0, 1, 1, 1, 1)).reshape(-1, Cout)
    out = ops.conv3x3_padded(xp, pack.pack_conv3x3(w), b, n_img, H, Wd, rowbias=rb, residual=res.contiguous())
    o = out.reshape(n_img, H + 2, Wd + 2, Cout)
    assert o[:, 0].abs().max() == 0 and o[:, -1].abs().max() == 0 and o[:, :, 0].abs().max() == 0 and o[:, :, -1].abs().max() == 0
    ref = F.conv2d(x.float(), bf(w).float(), b, padding=1) + rb[:, :, None, None]
    ref = ref.permute(0, 2, 3, 1) + res.reshape(n_img, H + 2, Wd + 2, Cout)[:, 1:-1, 1:-1].float()
    assert rel_l2(o[:, 1:-1, 1:-1], ref) < 5e-3, rel_l2(o[:, 1:-1, 1:-1], ref)


# ---------------------------------------------------------------------------------------------------
# norms at ds1
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("Cc,dense", [(960, True), (320, False)])
def test_groupnorm_ds1_bench_shape(ops, Cc, dense):
    """out9's in_layers GroupNorm over the virtual 640 + 320 concat (30-channel groups straddle the seam)
    with SiLU and the Plucker modulation, and the plain GroupNorm + SiLU of out_layers, at 42 x 72^2."""
    g = torch.Generator(device="cuda").manual_seed(Cc)
    n_img, H, Wd = N42, HW72, HW72
    x = bf(torch.randn(n_img, H, Wd, Cc, device="cuda", generator=g) * 1.5 + 0.3)
    gamma = torch.randn(Cc, device="cuda", generator=g)
    beta = torch.randn(Cc, device="cuda", generator=g)
    xin = F.pad(x, (0, 0, 1, 1, 1, 1)).reshape(-1, Cc).contiguous()
    dmap = dw = db = None
    if dense:
        dmap = torch.randn(n_img, H * Wd, 6, device="cuda", generator=g)
        dw = torch.randn(2 * Cc, 6, device="cuda", generator=g) * 0.3
        db = torch.randn(2 * Cc, device="cuda", generator=g) * 0.3
    out = ops.groupnorm(xin, gamma, beta, n_img, H, Wd, eps=1e-5, silu=True, dense=dmap, dense_w=dw, dense_b=db)
    o = out.reshape(n_img, H + 2, Wd + 2, Cc)
    assert o[:, 0].abs().max() == 0 and o[:, -1].abs().max() == 0 and o[:, :, 0].abs().max() == 0 and o[:, :, -1].abs().max() == 0
    err = []
    for n0 in range(0, n_img, 6):  # the fp32 reference in slices of 6 images (3.6 GB per temporary at C = 960)
        xs = x[n0:n0 + 6].permute(0, 3, 1, 2).float()
        ref = F.silu(F.group_norm(xs, 32, gamma, beta, 1e-5))
        if dense:
            d = torch.einsum("npk,ck->ncp", dmap[n0:n0 + 6], dw).reshape(-1, 2 * Cc, H, Wd) + db[None, :, None, None]
            sc, sh = d.chunk(2, dim=1)
            ref = ref * (1 + sc) + sh
        err.append(rel_l2(o[n0:n0 + 6, 1:-1, 1:-1].permute(0, 3, 1, 2), ref))
    assert max(err) < 4e-3, err


def test_layernorm_ds1_bench_shape(ops):
    g = torch.Generator(device="cuda").manual_seed(4)
    x = bf(torch.randn(M_DS1, 320, device="cuda", generator=g) * 2 + 0.5)
    gamma = torch.randn(320, device="cuda", generator=g)
    beta = torch.randn(320, device="cuda", generator=g)
    out = ops.layernorm(x, gamma, beta)
    ref = F.layer_norm(x.float(), (320,), gamma, beta, 1e-5)
    assert rel_l2(out, ref) < 4e-3
