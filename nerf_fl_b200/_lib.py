"""ctypes binding of libnerfw_b200.so (the C ABI declared in include/nerfw_b200.h).

There is no fallback: if the shared library is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# NWB_LIB_PATH: a differently-built libnerfw_b200.so (tools/build_variant.py: A/B runs of compile-time kernel options
# inside one GPU session); read once, at import
LIB_PATH = os.environ.get("NWB_LIB_PATH") or os.path.join(_HERE, "libnerfw_b200.so")

NWB_COMP_TRANSIENT, NWB_COMP_WHITE_BACK, NWB_COMP_TEST_TIME, NWB_COMP_COARSE, NWB_COMP_AMAX_CLEARED = 1, 2, 4, 8, 16
NWB_MLP_SIGMA_ONLY, NWB_MLP_TRANSIENT, NWB_MLP_SAVE, NWB_MLP_DZ_AMAX_GIVEN = 1, 2, 4, 8

_f32p = C.c_void_p   # device pointers travel as integers
_vp = C.c_void_p


class CompositeOut(C.Structure):
    _fields_ = [(n, _f32p) for n in (
        "weights", "opacity", "rgb", "depth", "transient_sigmas", "beta", "rgb_static",
        "rgb_transient", "tt_rgb_static", "tt_depth_static", "tt_rgb_transient",
        "tt_depth_transient", "clear_word")]


class CompositeGrads(C.Structure):
    _fields_ = [(n, _f32p) for n in (
        "d_weights", "d_opacity", "d_rgb", "d_depth", "d_transient_sigmas", "d_beta",
        "d_rgb_static", "d_rgb_transient")]


class LossSeed(C.Structure):
    """nwb_loss_seed: NerfWLoss cotangents formed inside the compositing backward (nwb_composite_bwd_seeded)."""
    _fields_ = [("rgb", _f32p), ("beta", _f32p), ("targets", _f32p), ("g_terms", _f32p * 5), ("coef", C.c_float),
                ("lambda_u", C.c_float), ("n_rays", C.c_int), ("n_ts", C.c_int64), ("fine", C.c_int)]


class FieldDesc(C.Structure):
    _fields_ = [("in_xyz", C.c_int), ("in_dir", C.c_int), ("in_a", C.c_int), ("in_t", C.c_int)]


_FIELD_SLOTS = ([("trunk_w", _f32p * 8), ("trunk_b", _f32p * 8)] +
                [(n, _f32p) for n in ("final_w", "final_b", "dir_w", "dir_b", "sigma_w", "sigma_b",
                                      "rgb_w", "rgb_b")] +
                [("tenc_w", _f32p * 4), ("tenc_b", _f32p * 4)] +
                [(n, _f32p) for n in ("tsigma_w", "tsigma_b", "trgb_w", "trgb_b", "tbeta_w",
                                      "tbeta_b")])


class FieldParams(C.Structure):
    _fields_ = _FIELD_SLOTS


class FieldGrads(C.Structure):
    _fields_ = _FIELD_SLOTS


_lib = None


def lib() -> C.CDLL:
    """Load (once) and return the shared library; raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -m nerf_fl_b200.build` "
            "(nvcc, sm_100a).  There is no CPU or PyTorch fallback.")
    L = C.CDLL(LIB_PATH)
    L.nwb_abi_version.restype = C.c_int
    L.nwb_last_error.restype = C.c_char_p
    i, u32, u64, f, sz = C.c_int, C.c_uint32, C.c_uint64, C.c_float, C.c_size_t
    protos = {
        "nwb_sample_coarse": [_vp, i, i, i, i, f, _vp, u64, _vp, _vp],
        "nwb_posenc": [_vp, C.c_int64, i, _vp, _vp],
        "nwb_sample_pdf_merge": [_vp, _vp, i, i, i, i, _vp, u64, _vp, _vp, _vp, _vp, _vp],
        "nwb_composite_fwd": [_vp, i, _vp, _vp, f, i, i, u32, f, C.POINTER(CompositeOut), _vp],
        "nwb_composite_bwd": [_vp, i, _vp, _vp, f, i, i, u32, C.POINTER(CompositeGrads), _vp, _vp],
        "nwb_composite_bwd_seeded": [_vp, i, _vp, _vp, f, i, i, u32, C.POINTER(CompositeGrads), C.POINTER(LossSeed), _vp,
                                     _vp, _vp],
        "nwb_selftest_umma": [_vp, _vp, i, i, i, C.POINTER(u32), _vp, _vp],
    }
    for name, args in protos.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = C.c_int
    _lib = L
    return L


def bind(name, argtypes, restype=C.c_int):
    """Late-bound prototype registration for entry points added by other modules."""
    fn = getattr(lib(), name)
    fn.argtypes = argtypes
    fn.restype = restype
    return fn


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().nwb_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"libnerfw_b200 {what} failed (code {rc}): {msg}")


def ptr(t: torch.Tensor | None):
    """Device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("nerf_fl_b200 kernels need CUDA tensors; there is no CPU path")
    if not t.is_contiguous():
        raise RuntimeError("nerf_fl_b200 kernels need contiguous tensors")
    return t.data_ptr()


def stream_ptr(device=None):
    return torch.cuda.current_stream(device).cuda_stream
