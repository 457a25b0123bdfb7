"""Mirror of ``kaolin.metrics.render`` (kaolin/metrics/render.py:18-40): ``mask_iou``, the loss that consumes
``soft_mask`` in the DIB-R fitting loop (examples/tutorial/ian_dibr.py:262-266), backed by libdibr_b200.so.

The reference runs eight torch kernels forward (mul, add, sub, two sums, add, div, mean) and their autograd
backward; here: one reduction pass over both masks (deterministic two-stage sum), one finishing kernel, and one
elementwise backward that writes ``grad_soft_mask`` analytically."""
import torch
from torch.autograd import Function

from .. import _lib

__all__ = ["mask_iou"]


class MaskIouCuda(Function):
    """forward(ctx, lhs (B,P), rhs (B,P)) -> scalar loss; backward -> (grad_lhs, grad_rhs)."""
    @staticmethod
    def forward(ctx, lhs, rhs):
        fn = "mask_iou"
        dev = lhs.device
        suf = _lib.suffix(lhs.dtype, fn)
        b, p = lhs.shape
        loss = torch.empty((), dtype=lhs.dtype, device=dev)
        sums = torch.empty((b, 2), dtype=lhs.dtype, device=dev)
        lib = _lib.load()
        with torch.cuda.device(dev):
            nbytes = lib.dibr_mask_iou_workspace_bytes(b, p, lhs.element_size())
            ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            code = getattr(lib, "dibr_mask_iou_fwd_" + suf)(
                _lib.ptr(lhs), _lib.ptr(rhs), b, p, _lib.ptr(loss), _lib.ptr(sums), _lib.ptr(ws), nbytes,
                _lib.current_stream(dev))
        _lib.check(code, fn)
        ctx.save_for_backward(lhs, rhs, sums)
        return loss

    @staticmethod
    def backward(ctx, grad_loss):
        fn = "mask_iou_backward"
        lhs, rhs, sums = ctx.saved_tensors
        dev = lhs.device
        suf = _lib.suffix(lhs.dtype, fn)
        g_lhs = torch.empty_like(lhs) if ctx.needs_input_grad[0] else None
        g_rhs = torch.empty_like(rhs) if ctx.needs_input_grad[1] else None
        b, p = lhs.shape
        with torch.cuda.device(dev):
            code = getattr(_lib.load(), "dibr_mask_iou_bwd_" + suf)(
                _lib.ptr(grad_loss.contiguous()), _lib.ptr(lhs), _lib.ptr(rhs), _lib.ptr(sums), b, p,
                _lib.ptr(g_lhs), _lib.ptr(g_rhs), _lib.current_stream(dev))
        _lib.check(code, fn)
        return g_lhs, g_rhs


def mask_iou(lhs_mask, rhs_mask):
    r"""Compute the Intersection over Union of two segmentation masks (reference:
    kaolin/metrics/render.py:18-40 — same arguments, same scalar return value).

    Args:
        lhs_mask (torch.FloatTensor): of shape :math:`(\text{batch_size}, \text{height}, \text{width})`.
        rhs_mask (torch.FloatTensor): of shape :math:`(\text{batch_size}, \text{height}, \text{width})`.

    Returns:
        (torch.FloatTensor): The IoU loss, as a torch scalar.
    """
    batch_size, height, width = lhs_mask.shape
    assert rhs_mask.shape == lhs_mask.shape
    if not lhs_mask.is_cuda or rhs_mask.device != lhs_mask.device:
        raise RuntimeError("mask_iou: both masks must be CUDA tensors on the same device "
                           "(libdibr_b200.so has no CPU path)")
    if rhs_mask.dtype != lhs_mask.dtype:
        rhs_mask = rhs_mask.to(lhs_mask.dtype)
    return MaskIouCuda.apply(lhs_mask.reshape(batch_size, -1).contiguous(),
                             rhs_mask.reshape(batch_size, -1).contiguous())


class ImageL1Cuda(Function):
    """forward(ctx, a, b) -> mean |a - b| (scalar); backward -> (grad_a, grad_b)."""
    @staticmethod
    def forward(ctx, a, b):
        fn = "image_l1"
        dev = a.device
        suf = _lib.suffix(a.dtype, fn)
        n = a.numel()
        loss = torch.empty((), dtype=a.dtype, device=dev)
        lib = _lib.load()
        with torch.cuda.device(dev):
            nbytes = lib.dibr_image_l1_workspace_bytes(n, a.element_size())
            ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            code = getattr(lib, "dibr_image_l1_fwd_" + suf)(_lib.ptr(a), _lib.ptr(b), n, _lib.ptr(loss), _lib.ptr(ws),
                                                            nbytes, _lib.current_stream(dev))
        _lib.check(code, fn)
        ctx.save_for_backward(a, b)
        return loss

    @staticmethod
    def backward(ctx, grad_loss):
        fn = "image_l1_backward"
        a, b = ctx.saved_tensors
        dev = a.device
        suf = _lib.suffix(a.dtype, fn)
        g_a = torch.empty_like(a) if ctx.needs_input_grad[0] else None
        g_b = torch.empty_like(b) if ctx.needs_input_grad[1] else None
        with torch.cuda.device(dev):
            code = getattr(_lib.load(), "dibr_image_l1_bwd_" + suf)(
                _lib.ptr(grad_loss.contiguous()), _lib.ptr(a), _lib.ptr(b), a.numel(), _lib.ptr(g_a), _lib.ptr(g_b),
                _lib.current_stream(dev))
        _lib.check(code, fn)
        return g_a, g_b


def image_l1(image, gt_image):
    r"""``torch.mean(torch.abs(image - gt_image))`` — the image loss of the DIB-R fitting loop
    (examples/tutorial/ian_dibr.py:260; not a function of the reference's ``kaolin.metrics.render``, which composes it
    from torch ops) as one deterministic reduction forward and one elementwise kernel backward.

    Args:
        image (torch.FloatTensor): rendered image, any shape.
        gt_image (torch.FloatTensor): target of the same shape.

    Returns:
        (torch.FloatTensor): the mean absolute difference, as a torch scalar.
    """
    assert gt_image.shape == image.shape
    if not image.is_cuda or gt_image.device != image.device:
        raise RuntimeError("image_l1: both images must be CUDA tensors on the same device "
                           "(libdibr_b200.so has no CPU path)")
    if gt_image.dtype != image.dtype:
        gt_image = gt_image.to(image.dtype)
    return ImageL1Cuda.apply(image.contiguous(), gt_image.contiguous())
