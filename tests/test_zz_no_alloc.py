"""dibr_mask_iou_fwd_noalloc_* — the allocation-free variant of the mask_iou forward (include/dibr_b200.h; no
counterpart in the reference).  Its scratch is a file-scope
`static __device__` array inside the library instead of a caller workspace.

Called through the C ABI (ctypes) and compared with the existing entry point dibr_mask_iou_fwd_* on the same
inputs, on two calls in a row (the second with other sizes, so that it runs over the first call's leftovers in the
shared array), fp32 and fp64.  Both entry points run the same deterministic two-stage reduction, so the results are
expected to be bit-equal; the assertion uses the tolerance of tests/test_texture_losses.py (1e-6 / 1e-12 relative).

This file sorts last on purpose (a failure here must not hide other tests) and passes both
alone in a fresh process and after the rest of the suite."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = {torch.float32: 1e-6, torch.float64: 1e-12}


def _close(a, b, tol):
    a, b = a.double().cpu(), b.double().cpu()
    return bool(((a - b).abs() <= tol * b.abs().clamp(min=1.0)).all())


def _existing(lib, suf, lhs, rhs, stream):
    b, p = lhs.shape
    loss = torch.full((), float("nan"), dtype=lhs.dtype, device=lhs.device)
    sums = torch.full((b, 2), float("nan"), dtype=lhs.dtype, device=lhs.device)
    n = lib.dibr_mask_iou_workspace_bytes(b, p, lhs.element_size())
    ws = torch.empty(max(int(n), 256), dtype=torch.uint8, device=lhs.device)
    code = getattr(lib, "dibr_mask_iou_fwd_" + suf)(lhs.data_ptr(), rhs.data_ptr(), b, p, loss.data_ptr(),
                                                    sums.data_ptr(), ws.data_ptr(), n, stream)
    assert code == 0, lib.dibr_error_string(code)
    return loss, sums


def _noalloc(lib, suf, lhs, rhs, stream):
    """Outputs are allocated by the caller BEFORE the call; the call itself must not allocate."""
    b, p = lhs.shape
    loss = torch.full((), float("nan"), dtype=lhs.dtype, device=lhs.device)
    sums = torch.full((b, 2), float("nan"), dtype=lhs.dtype, device=lhs.device)
    torch.cuda.synchronize()
    before = (torch.cuda.memory_allocated(), torch.cuda.memory_reserved())
    code = getattr(lib, "dibr_mask_iou_fwd_noalloc_" + suf)(lhs.data_ptr(), rhs.data_ptr(), b, p, loss.data_ptr(),
                                                            sums.data_ptr(), stream)
    assert code == 0, lib.dibr_error_string(code)
    torch.cuda.synchronize()
    assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == before, "the call went through torch's allocator"
    return loss, sums


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_noalloc_mask_iou_matches_existing_entry_point_twice(dtype):
    from kaolin_b200 import _lib
    lib = _lib.load()
    suf = _lib.suffix(dtype, "mask_iou")
    dev = torch.device("cuda", 0)
    assert lib.dibr_noalloc_scratch_bytes() == 1 << 20
    with torch.cuda.device(dev):
        stream = _lib.current_stream(dev)
        g = torch.Generator().manual_seed(2024)
        # two calls in a row on the shared scratch: a large one (7 views x 300 x 333: 13 chunks per view, ragged
        # last chunk), then a smaller one that reuses the front of the array over the first call's leftovers
        for b, p in ((7, 300 * 333), (3, 64 * 50), (7, 300 * 333)):
            lhs = torch.rand(b, p, generator=g, dtype=torch.float64).to(dev, dtype)
            rhs = (torch.rand(b, p, generator=g) > 0.4).to(dev, dtype)
            want_loss, want_sums = _existing(lib, suf, lhs, rhs, stream)
            got_loss, got_sums = _noalloc(lib, suf, lhs, rhs, stream)
            assert torch.isfinite(got_sums).all() and torch.isfinite(got_loss)
            assert _close(got_sums, want_sums, TOL[dtype]), (got_sums, want_sums)
            assert _close(got_loss, want_loss, TOL[dtype]), (float(got_loss), float(want_loss))
            # and against the definition (kaolin/metrics/render.py:18-40) in float64
            l64, r64 = lhs.double(), rhs.double()
            up, down = (l64 * r64).sum(1), (l64 + r64 - l64 * r64).sum(1)
            ref = 1.0 - (up / (down + 1e-10)).mean()
            assert abs(float(got_loss) - float(ref)) <= (1e-5 if dtype == torch.float32 else 1e-12)


def test_noalloc_mask_iou_rejects_sizes_beyond_the_built_in_scratch():
    from kaolin_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    with torch.cuda.device(dev):
        stream = _lib.current_stream(dev)
        limit = lib.dibr_noalloc_scratch_bytes()
        b, p = 4096, 8192 * 33                       # 4096 x 33 chunks x 8 bytes > 1 MiB
        assert lib.dibr_mask_iou_workspace_bytes(b, p, 4) > limit
        out = torch.zeros(b * 2 + 1, device=dev)
        null = ctypes.c_void_p(0)
        # rejected on the sizes alone, before any pointer is used
        code = lib.dibr_mask_iou_fwd_noalloc_f32(null, null, b, p, out.data_ptr(), out.data_ptr() + 4, stream)
        assert code == 5 and b"scratch" in lib.dibr_error_string(code)
        # the largest accepted size class still works: B * chunks * 8 bytes == 1 MiB exactly
        b, p = 2048, 8192 * 64 - 5
        assert lib.dibr_mask_iou_workspace_bytes(b, p, 4) == limit
        lhs = torch.full((b, p), 0.5, device=dev)
        rhs = torch.ones((b, p), device=dev)
        loss, sums = _noalloc(lib, "f32", lhs, rhs, stream)
        # up = 0.5 P, down = 1.0 P -> IoU 0.5 for every view
        assert abs(float(loss) - 0.5) < 1e-5
        torch.cuda.synchronize()
