"""tcgen05 multi-tap GEMM vs torch (fp64) on the same bf16 hi/lo operands — through the C ABI."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _hi_dtype():
    """Plane layout of the loaded library (include/ramp_b200.h: ramp_plane_format): fp16 + fp16 (2) or bf16 + bf16 (0)."""
    from ramp_b200._lib import lib
    return torch.float16 if int(lib().ramp_plane_format()) == 2 else torch.bfloat16


def _split(x):
    hi = x.clamp(-65504.0, 65504.0).to(_hi_dtype())
    lo = (x - hi.float()).to(_hi_dtype())
    return hi.contiguous(), lo.contiguous()


def _gemm(a, w, taps, L_out, a_pstride=1, bias=None, bias_tab=None, row2scene=None, residual=None, precision=3,
          o_L=None, o_stride=1, o_off=0, out_rows=None, K=None, want_split=False):
    """a: [n, L_in, ldA] fp32 cuda; w: [n_wtaps, Nout, Kw]."""
    from ramp_b200._lib import lib, check, ptr, stream_ptr
    n, L_in, ldA = a.shape
    n_wt, Nout, Kw = w.shape
    K = K or Kw
    ah, al = _split(a)
    wh, wl = _split(w)
    o_L = o_L or L_out
    out_rows = out_rows or n * o_L
    # guard bands of 160 rows around every output (compute-sanitizer is closed on this pool): the kernel masks / the TMA
    # unit clips the rows of a partial last tile, so the bands must come back untouched
    G, SENT = 160, 12288.0

    def banded(dtype):
        buf = torch.full((G + out_rows + G, Nout), SENT, device=a.device, dtype=dtype)
        buf[G:G + out_rows] = 0
        return buf, buf[G:G + out_rows]
    out_buf, out = banded(torch.float32)
    oh_buf, oh = banded(_hi_dtype()) if want_split else (None, None)
    ol_buf, ol = banded(_hi_dtype()) if want_split else (None, None)
    tp = (ctypes.c_int * (3 * len(taps)))(*[v for t in taps for v in t])
    check(lib().ramp_gemm(ptr(ah), ptr(al), n, L_in, ldA, K, a_pstride, L_out, len(taps), tp, ptr(wh), ptr(wl), Nout, Kw,
                          n_wt, ptr(bias), ptr(bias_tab), ptr(row2scene), ptr(residual), Nout, ptr(out), Nout,
                          ptr(oh), ptr(ol), Nout, o_L, o_stride, o_off, precision, stream_ptr()), "ramp_gemm")
    torch.cuda.synchronize()
    for buf in (out_buf, oh_buf, ol_buf):
        if buf is not None:
            assert bool((buf[:G] == SENT).all()) and bool((buf[G + out_rows:] == SENT).all()), "write outside the output rows"
    ar = ah.double() + (al.double() if precision == 3 else 0)
    wr = wh.double() + (wl.double() if precision == 3 else 0)
    return out, (oh, ol), ar, wr


@pytest.mark.parametrize("n,L,K,Nout", [(2, 64, 64, 32), (16, 8, 256, 256), (4, 32, 256, 768), (3, 64, 2048, 256),
                                        (5, 16, 32, 2048), (2, 64, 16, 32), (2, 64, 32, 16), (33, 8, 768, 256)])
@pytest.mark.parametrize("precision", [3, 1])
def test_linear(cuda, n, L, K, Nout, precision):
    g = torch.Generator(device="cpu").manual_seed(n * 131 + K)
    a = torch.randn(n, L, K, generator=g).to(cuda)
    w = (torch.randn(1, Nout, K, generator=g) / K ** 0.5).to(cuda)
    bias = torch.randn(Nout, generator=g).to(cuda)
    res = torch.randn(n * L, Nout, generator=g).to(cuda)
    out, _, ar, wr = _gemm(a, w, [(0, 0, 0)], L, bias=bias, residual=res, precision=precision)
    ref = ar.reshape(-1, K) @ wr[0].T + bias.double() + res.double()
    err = (out.double() - ref).abs().max().item()
    # dropped lo*lo term and fp32 accumulation
    assert err < (2e-5 if precision == 3 else 2e-4) * max(1.0, ref.abs().max().item()), err


def test_split_output_and_scene_bias(cuda):
    g = torch.Generator(device="cpu").manual_seed(7)
    n, L, K, Nout = 6, 32, 256, 256
    a = torch.randn(n, L, K, generator=g).to(cuda)
    w = (torch.randn(1, Nout, K, generator=g) / 16).to(cuda)
    tab = torch.randn(3, Nout, generator=g).to(cuda)
    r2s = torch.tensor([0, 1, 2, 2, 1, 0], dtype=torch.int32, device=cuda)
    out, (oh, ol), ar, wr = _gemm(a, w, [(0, 0, 0)], L, bias_tab=tab, row2scene=r2s, want_split=True)
    ref = (ar.reshape(-1, K) @ wr[0].T).reshape(n, L, Nout) + tab.double()[r2s.long()][:, None, :]
    assert (out.double().reshape(n, L, Nout) - ref).abs().max().item() < 5e-5
    assert (oh.float() + ol.float() - out).abs().max().item() < 1e-4


@pytest.mark.parametrize("n,L,Cin,Cout", [(2, 64, 32, 32), (4, 32, 64, 64), (16, 8, 512, 128), (3, 16, 128, 128), (2, 64, 16, 32)])
def test_conv5(cuda, n, L, Cin, Cout):
    g = torch.Generator(device="cpu").manual_seed(L + Cin)
    a = torch.randn(n, L, Cin, generator=g).to(cuda)
    wt = (torch.randn(Cout, Cin, 5, generator=g) / (5 * Cin) ** 0.5)
    w = wt.permute(2, 0, 1).contiguous().to(cuda)      # [tap, Cout, Cin]
    bias = torch.randn(Cout, generator=g).to(cuda)
    out, _, ar, wr = _gemm(a, w, [(0, k - 2, k) for k in range(5)], L, bias=bias)
    ref = F.conv1d(ar.transpose(1, 2), wr.permute(1, 2, 0), bias.double(), padding=2).transpose(1, 2)
    assert (out.double().reshape(n, L, Cout) - ref).abs().max().item() < 5e-5


def test_stride2_down(cuda):
    """Conv1d k3 s2 p1 (layers.py:262-268) through the parity view of the input."""
    g = torch.Generator(device="cpu").manual_seed(3)
    n, L, C = 4, 32, 64
    a = torch.randn(n, L, C, generator=g).to(cuda)
    wt = torch.randn(C, C, 3, generator=g) / (3 * C) ** 0.5
    w = wt.permute(2, 0, 1).contiguous().to(cuda)
    out, _, ar, wr = _gemm(a, w, [(1, -1, 0), (0, 0, 1), (1, 0, 2)], L // 2, a_pstride=2)
    ref = F.conv1d(ar.transpose(1, 2), wr.permute(1, 2, 0), stride=2, padding=1).transpose(1, 2)
    assert (out.double().reshape(n, L // 2, C) - ref).abs().max().item() < 5e-5


def test_transposed_up(cuda):
    """ConvTranspose1d k4 s2 p1 (layers.py:271-277) as two parity GEMMs with interleaved output rows."""
    g = torch.Generator(device="cpu").manual_seed(4)
    n, L, C = 4, 16, 128
    a = torch.randn(n, L, C, generator=g).to(cuda)
    wt = torch.randn(C, C, 4, generator=g) / (2 * C) ** 0.5          # [Cin, Cout, k]
    w = wt.permute(2, 1, 0).contiguous().to(cuda)                     # [k, Cout, Cin]
    ev, _, ar, wr = _gemm(a, w, [(0, 0, 1), (0, -1, 3)], L, o_L=2 * L, o_stride=2, o_off=0, out_rows=n * 2 * L)
    od, _, _, _ = _gemm(a, w, [(0, 1, 0), (0, 0, 2)], L, o_L=2 * L, o_stride=2, o_off=1, out_rows=n * 2 * L)
    out = (ev + od).double().reshape(n, 2 * L, C)
    ref = F.conv_transpose1d(ar.transpose(1, 2), wr.permute(2, 1, 0), stride=2, padding=1).transpose(1, 2)
    assert (out - ref).abs().max().item() < 5e-5


@pytest.mark.parametrize("n,L,K,Nout", [(600, 64, 256, 256), (601, 64, 256, 768), (2400, 8, 2048, 256), (1203, 16, 256, 2048)])
def test_linear_many_tiles(cuda, n, L, K, Nout):
    """Enough M tiles to fill the GPU (and an odd tile count): the shapes where RAMP_GEMM_PAIR=1 puts a 256-row tile on a CTA
    pair (tcgen05 cta_group::2, half of the weight rows staged per CTA); the same test runs in either mode."""
    g = torch.Generator(device="cpu").manual_seed(n + K)
    a = torch.randn(n, L, K, generator=g).to(cuda)
    w = (torch.randn(1, Nout, K, generator=g) / K ** 0.5).to(cuda)
    bias = torch.randn(Nout, generator=g).to(cuda)
    res = torch.randn(n * L, Nout, generator=g).to(cuda)
    out, (oh, ol), ar, wr = _gemm(a, w, [(0, 0, 0)], L, bias=bias, residual=res, want_split=True)
    # fp64 reference in row blocks (the full product is 38400 x 2048 x 256)
    err = 0.0
    for s in range(0, n * L, 8192):
        r = ar.reshape(-1, K)[s:s + 8192] @ wr[0].T + bias.double() + res[s:s + 8192].double()
        err = max(err, (out[s:s + 8192].double() - r).abs().max().item() / max(1.0, r.abs().max().item()))
    assert err < 2e-5, err
    assert (oh.float() + ol.float() - out).abs().max().item() < 1e-4
