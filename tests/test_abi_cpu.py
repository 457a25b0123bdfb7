"""CPU checks of the drop-in boundary: the C-ABI library loads and exports every function include/dibr_b200.h
declares (no compute calls — there is no GPU here), the pure-host entry points answer, and the Python mirror
keeps the reference's argument / error conventions (kaolin/render/mesh/rasterization.py:390-506, dibr.py:75-209,
kaolin/csrc/render/mesh/rasterization.cpp:60-85)."""
import ctypes
import inspect
import os
import re
import warnings

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from kaolin_b200 import _lib
    _lib.build()
    return _lib.load()


def declared_functions():
    text = open(os.path.join(ROOT, "include", "dibr_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dibr_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(lib):
    from kaolin_b200 import _lib
    names = declared_functions()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"libdibr_b200.so does not export {n}"
    assert set(_lib.EXPORTED_SYMBOLS) == set(names), set(_lib.EXPORTED_SYMBOLS) ^ set(names)


def test_host_only_entry_points(lib):
    assert lib.dibr_abi_version() == 8
    assert lib.dibr_error_string(0) == b"ok"
    for code in (1, 2, 3, 4, 5, 99):
        assert len(lib.dibr_error_string(code)) > 0
    n = lib.dibr_workspace_bytes(16, 16 * 20000, 512, 512)
    assert n > 16 * 20000 * 8 and n % 256 == 0
    assert lib.dibr_workspace_bytes(-1, 10, 4, 4) == 0
    # a pure function of the sizes, monotone in the face count
    assert lib.dibr_workspace_bytes(2, 1000, 64, 64) == lib.dibr_workspace_bytes(2, 1000, 64, 64)
    assert lib.dibr_workspace_bytes(2, 2000, 64, 64) > lib.dibr_workspace_bytes(2, 1000, 64, 64)
    lib.dibr_reset_kernel_launch_count()
    assert lib.dibr_kernel_launch_count() == 0


def test_invalid_arguments_are_rejected_before_any_cuda_call(lib):
    null = ctypes.c_void_p(0)
    # negative sizes / missing workspace -> error codes, no exception, no CUDA call needed
    code = lib.dibr_rasterize_fwd_fused_f32(null, null, null, null, -1, 4, 3, 8, 8, 1000.0, 1e-8, null, null, null,
                                            null, 0, null)
    assert code == 1
    code = lib.dibr_rasterize_fwd_fused_f32(null, null, null, null, 1, 4, 3, 40000, 8, 1000.0, 1e-8, null, null,
                                            null, null, 0, null)
    assert code == 4
    code = lib.dibr_soft_mask_fwd_fused_f32(null, null, 1, 4, 8, 8, 7000.0, 0.02, -3, 1000.0, null, null, null,
                                            null, null, 0, 0, null)
    assert code == 1
    assert lib.dibr_profile_next_launch(3, null, null) == 1
    assert lib.dibr_profile_next_launch(-1, null, null) == 0


def test_public_api_signatures_match_reference():
    import kaolin_b200 as kal
    m = kal.render.mesh
    assert list(inspect.signature(m.rasterize).parameters) == [
        "height", "width", "face_vertices_z", "face_vertices_image", "face_features", "valid_faces", "multiplier",
        "eps", "backend"]
    assert inspect.signature(m.rasterize).parameters["backend"].default == "cuda"
    sig = inspect.signature(m.dibr_soft_mask)
    assert [(k, v.default) for k, v in sig.parameters.items()][2:] == [
        ("sigmainv", 7000), ("boxlen", 0.02), ("knum", 30), ("multiplier", 1000.)]
    sig = inspect.signature(m.dibr_rasterization)
    assert list(sig.parameters) == ["height", "width", "face_vertices_z", "face_vertices_image", "face_features",
                                    "face_normals_z", "sigmainv", "boxlen", "knum", "multiplier", "eps",
                                    "rast_backend"]
    for name in ("packed_rasterize_forward_cuda", "rasterize_backward_cuda", "dibr_soft_mask_forward_cuda",
                 "dibr_soft_mask_backward_cuda"):
        assert callable(getattr(kal._C.render.mesh, name))
    from kaolin_b200.render.mesh import rasterization, dibr
    assert rasterization.__all__ == ["rasterize"] and dibr.__all__ == ["dibr_soft_mask", "dibr_rasterization"]
    for cls in (rasterization.RasterizeCuda, rasterization.NvdiffRasterizeFwdCudaBwd, dibr.DibrSoftMaskCuda):
        assert issubclass(cls, torch.autograd.Function)


def test_error_conventions_without_gpu():
    import kaolin_b200 as kal
    z = torch.zeros(1, 2, 3)
    img = torch.zeros(1, 2, 3, 2)
    feat = torch.zeros(1, 2, 3, 4)
    with pytest.raises(ValueError, match="not a valid backend"):
        kal.render.mesh.rasterize(4, 4, z, img, feat, backend="opengl")
    with pytest.raises(ValueError, match="nvdiffrast must be installed"):      # unavailable on this box
        kal.render.mesh.rasterize(4, 4, z, img, feat, backend="nvdiffrast")
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        with pytest.raises(ValueError):
            kal.render.mesh.rasterize(4, 4, z, img, feat, multiplier=100, eps=1e-6, backend="nvdiffrast")
        assert any("multiplier is ignored" in str(x.message) for x in w)
        assert any("eps is ignored" in str(x.message) for x in w)
    # CPU tensors: the reference's at::checkAllSameGPU -> RuntimeError; there is NO CPU fallback
    with pytest.raises(RuntimeError, match="CUDA"):
        kal.render.mesh.rasterize(4, 4, z, img, feat)
    with pytest.raises(RuntimeError, match="CUDA"):
        kal.render.mesh.dibr_soft_mask(img, torch.zeros(1, 4, 4, dtype=torch.long))
    with pytest.raises(RuntimeError, match="CUDA"):
        kal._C.render.mesh.packed_rasterize_forward_cuda(4, 4, z[0], img[0], torch.zeros(2, 4), feat[0],
                                                         torch.tensor([0, 2]), 1000., 1e-8)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "kaolin_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text and "oracle/" not in text, f


def test_round2_entry_points_fail_loudly_without_gpu():
    """The additions of round 2 (from-mesh entry, image L1, per-element Adam) have no CPU path either."""
    import kaolin_b200 as kal
    from kaolin_b200.metrics.render import image_l1
    from kaolin_b200.multiview import PerElementAdam
    v = torch.rand(1, 12, 3)
    f = torch.tensor([[0, 1, 2], [2, 3, 4]], dtype=torch.long)
    feats = torch.rand(2, 2, 3, 3)
    proj = torch.tensor([2.4, 2.4, -1.0])
    rot = torch.eye(3).repeat(2, 1, 1)
    trans = torch.zeros(2, 3)
    with pytest.raises(RuntimeError, match="CUDA"):
        kal.render.mesh.dibr_rasterization_from_mesh(8, 8, v, f, feats, proj, rot, trans)
    with pytest.raises(AssertionError):                       # prepare_vertices' own assertion (utils.py:160-166)
        kal.render.mesh.dibr_rasterization_from_mesh(8, 8, v, f, feats, proj)
    with pytest.raises(RuntimeError, match="CUDA"):
        image_l1(torch.zeros(2, 4, 4, 3), torch.zeros(2, 4, 4, 3))
    p = torch.zeros(4, requires_grad=True)
    p.grad = torch.ones(4)
    with pytest.raises(RuntimeError, match="CUDA"):
        PerElementAdam([p]).step()
    # the reference's module surface is unchanged; the additions are exported next to it
    from kaolin_b200.render.mesh import dibr
    assert dibr.__all__ == ["dibr_soft_mask", "dibr_rasterization"]
    assert callable(kal.render.mesh.dibr_rasterization_from_mesh)
