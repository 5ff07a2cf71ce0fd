"""seva_groupnorm_stats_static: the allocation-free variant of the GroupNorm statistics kernel (include/seva_b200.h), called
through the C ABI and checked against the existing entry point seva_groupnorm_stats on the same inputs.

The variant keeps its scratch (per-block partials + tickets) in a device array inside the library, which every call shares
and which keeps the previous call's contents: the test therefore makes two calls in a row with DIFFERENT shapes (the second
one finds the first one's partials in the array) and then repeats the first.  Named test_zz_* so that pytest collects it last.
"""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "stable-virtual-camera_b200")]

pytestmark = pytest.mark.gpu

MAX_FLOATS = 1 << 22   # SEVA_GN_STATS_STATIC_MAX_FLOATS
MAX_IMAGES = 1024      # SEVA_GN_STATS_STATIC_MAX_IMAGES


def _rel_l2(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm()).item()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _existing(lib, x, n_img, rpi, C, reverse):
    n = lib.seva_groupnorm_stats_floats(n_img, rpi, C)
    assert n > 0
    ws = torch.zeros(n, device="cuda", dtype=torch.float32)
    rc = lib.seva_groupnorm_stats(x.data_ptr(), x.stride(0), n_img, rpi, C, ws.data_ptr(), int(reverse), _stream())
    assert rc == 0
    return ws[: n_img * C * 2].clone()


def _static(lib, _lib, x, n_img, rpi, C, reverse, out):
    rc = lib.seva_groupnorm_stats_static(x.data_ptr(), x.stride(0), n_img, rpi, C, out.data_ptr(), int(reverse), _stream())
    assert rc == 0, _lib.last_error()
    return out


def test_static_scratch_stats_match_the_workspace_entry_point():
    from seva_b200 import _lib

    lib = _lib.load()
    g = torch.Generator(device="cuda").manual_seed(11)
    # (images, rows per image, channels): a padded 38^2 level, the step's largest GroupNorm shape (42 x 74^2 x 320) and a
    # ragged one (rows not a multiple of the block's share); the first shape is repeated at the end
    shapes = [(5, 38 * 38, 640), (42, 74 * 74, 320), (3, 1001, 1280), (5, 38 * 38, 640)]
    for i, (n_img, rpi, C) in enumerate(shapes):
        x = torch.randn(n_img * rpi, C, device="cuda", generator=g).to(torch.bfloat16)
        want = _existing(lib, x, n_img, rpi, C, reverse=i & 1)
        out = torch.full((n_img * C * 2,), float("nan"), device="cuda", dtype=torch.float32)
        got = _static(lib, _lib, x, n_img, rpi, C, i & 1, out)
        # same kernel, same launch shape, fixed summation order: the bits agree; the tolerance is the stats tests' own
        assert _rel_l2(got, want) < 1e-5
        assert torch.equal(got, want)
        # and again on the same input (the tickets were left at zero, the partials are overwritten)
        out2 = torch.empty_like(out)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        again = _static(lib, _lib, x, n_img, rpi, C, i & 1, out2)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before          # the call itself took no device memory
        assert torch.equal(again, want)
        ref = torch.stack([x.float().view(n_img, rpi, C).sum(1), (x.float() ** 2).view(n_img, rpi, C).sum(1)], -1)
        assert _rel_l2(got.view(n_img, C, 2), ref) < 1e-5


def test_static_scratch_rejects_problems_beyond_its_maximum():
    from seva_b200 import _lib

    lib = _lib.load()
    x = torch.zeros(64, 1280, device="cuda", dtype=torch.bfloat16)
    out = torch.zeros(8, device="cuda")
    # more images than tickets
    rc = lib.seva_groupnorm_stats_static(x.data_ptr(), x.stride(0), MAX_IMAGES + 1, 16, 64, out.data_ptr(), 0, _stream())
    assert rc == 1 and "static scratch" in _lib.last_error() and str(MAX_IMAGES) in _lib.last_error()
    # more per-block partials than the array holds; rejected before any launch, so the undersized x / out are never touched
    n_img, rpi, C = 8, 1000, 4096
    partials = lib.seva_groupnorm_stats_floats(n_img, rpi, C) - n_img * C * 2 - ((n_img + 3) & ~3)
    assert partials > MAX_FLOATS, "pick a larger problem: this one fits the static scratch"
    rc = lib.seva_groupnorm_stats_static(x.data_ptr(), 4096, n_img, rpi, C, out.data_ptr(), 0, _stream())
    assert rc == 1 and "static scratch" in _lib.last_error() and str(MAX_FLOATS) in _lib.last_error()
    torch.cuda.synchronize()
    # the entry point still works after a rejection
    xs = torch.randn(2 * 100, 64, device="cuda").to(torch.bfloat16)
    got = _static(lib, _lib, xs, 2, 100, 64, 0, torch.empty(2 * 64 * 2, device="cuda"))
    assert torch.equal(got, _existing(lib, xs, 2, 100, 64, 0))
