"""ctypes binding of ``libcm_b200.so`` (C ABI declared in ``include/cm_b200.h``).

PyTorch is used for device memory and streams only: every call below hands raw device pointers and the
current CUDA stream to the library.  There is no CPU or eager fallback -- a missing library, a missing
symbol or a non-sm_100 device raises.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcm_b200.so")

CM_OK, CM_ERR_ARG, CM_ERR_DEVICE, CM_ERR_LAUNCH, CM_ERR_WORKSPACE = 0, -1, -2, -3, -4
STATUS_LABEL_OVERFLOW, STATUS_NODE_OVERFLOW, STATUS_EDGE_OVERFLOW = 1, 2, 4
NODE_FEATURES = 15

_fp = C.c_void_p   # every device pointer crosses as void*


class GnnParams(C.Structure):
    _fields_ = [
        ("in_channels", C.c_int), ("hidden", C.c_int), ("heads", C.c_int), ("num_classes", C.c_int),
        ("conv1_lin_weight", _fp), ("conv1_att_src", _fp), ("conv1_att_dst", _fp), ("conv1_bias", _fp),
        ("conv_lin_weight", _fp * 3), ("conv_bias", _fp * 3),
        ("bn_weight", _fp * 4), ("bn_bias", _fp * 4), ("bn_mean", _fp * 4), ("bn_var", _fp * 4),
        ("bn_eps", C.c_float),
        ("fc_shared_weight", _fp), ("fc_shared_bias", _fp),
        ("head1_weight", _fp * 3), ("head1_bias", _fp * 3), ("head2_weight", _fp * 3), ("head2_bias", _fp * 3),
    ]


class GnnDropout(C.Structure):
    _fields_ = [("keep", _fp * 7), ("p_conv", C.c_float), ("p_fc", C.c_float)]


class GnnGrads(C.Structure):
    _fields_ = [
        ("conv1_lin_weight", _fp), ("conv1_att_src", _fp), ("conv1_att_dst", _fp), ("conv1_bias", _fp),
        ("conv_lin_weight", _fp * 3), ("conv_bias", _fp * 3),
        ("bn_weight", _fp * 4), ("bn_bias", _fp * 4),
        ("fc_shared_weight", _fp), ("fc_shared_bias", _fp),
        ("head1_weight", _fp * 3), ("head1_bias", _fp * 3), ("head2_weight", _fp * 3), ("head2_bias", _fp * 3),
    ]


class FusionParams(C.Structure):
    _fields_ = [
        ("rg_dim", C.c_int), ("kg_dim", C.c_int), ("hidden", C.c_int), ("heads", C.c_int),
        ("num_classes", C.c_int), ("fusion_type", C.c_int),
        ("rg_proj_w", _fp), ("rg_proj_b", _fp), ("kg_proj_w", _fp), ("kg_proj_b", _fp),
        ("r2k_in_w", _fp), ("r2k_in_b", _fp), ("r2k_out_w", _fp), ("r2k_out_b", _fp),
        ("k2r_in_w", _fp), ("k2r_in_b", _fp), ("k2r_out_w", _fp), ("k2r_out_b", _fp),
        ("ln_rg_w", _fp), ("ln_rg_b", _fp), ("ln_kg_w", _fp), ("ln_kg_b", _fp),
        ("ffn_rg_w1", _fp), ("ffn_rg_b1", _fp), ("ffn_rg_w2", _fp), ("ffn_rg_b2", _fp),
        ("ffn_kg_w1", _fp), ("ffn_kg_b1", _fp), ("ffn_kg_w2", _fp), ("ffn_kg_b2", _fp),
        ("fl_w1", _fp), ("fl_b1", _fp), ("fl_w2", _fp), ("fl_b2", _fp),
        ("late_w", _fp * 3), ("late_b", _fp * 3),
        ("head_w1", _fp * 4), ("head_b1", _fp * 4), ("head_w2", _fp * 4), ("head_b2", _fp * 4),
    ]


class RgLossConfig(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("mask_pos_weight", "instance_pos_weight", "edge_pos_weight", "task_mask", "task_instance", "task_edge")]


class ImageLossConfig(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("alpha", "gamma", "w_focal", "w_instance", "w_edge", "w_score")]


class AdamWConfig(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("lr", "beta1", "beta2", "eps", "weight_decay", "max_norm")]


class FusionDropout(C.Structure):
    _fields_ = [("keep", _fp * 9), ("p", C.c_float)]


FUSION_GRAD_FIELDS = ["rg_proj_w", "rg_proj_b", "kg_proj_w", "kg_proj_b", "r2k_in_w", "r2k_in_b", "r2k_out_w", "r2k_out_b",
                      "k2r_in_w", "k2r_in_b", "k2r_out_w", "k2r_out_b", "ln_rg_w", "ln_rg_b", "ln_kg_w", "ln_kg_b",
                      "ffn_rg_w1", "ffn_rg_b1", "ffn_rg_w2", "ffn_rg_b2", "ffn_kg_w1", "ffn_kg_b1", "ffn_kg_w2", "ffn_kg_b2",
                      "fl_w1", "fl_b1", "fl_w2", "fl_b2"]


class FusionGrads(C.Structure):
    _fields_ = [(n, _fp) for n in FUSION_GRAD_FIELDS] + [("head_w1", _fp * 4), ("head_b1", _fp * 4), ("head_w2", _fp * 4),
                                                         ("head_b2", _fp * 4), ("late_w", _fp * 3), ("late_b", _fp * 3)]


# name -> (restype, argtypes); the list is also what tests/test_abi.py checks against include/cm_b200.h
_i, _sz, _f, _i64 = C.c_int, C.c_size_t, C.c_float, C.c_int64
SIGNATURES = {
    "cm_last_error": (C.c_char_p, []),
    "cm_abi_version": (_i, []),
    "cm_device_check": (_i, []),
    "cm_timing_enable": (None, [_i]),
    "cm_debug_fusion_phase_stamps": (None, [_fp]),
    "cm_timing_report": (_i, [C.c_char_p, _i]),
    "cm_rag_workspace_bytes": (_sz, [_i, _i]),
    "cm_rag_build": (_i, [_fp, _i, _i, _i, _i, _fp, _sz, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _i, _i, _i, _fp, _fp]),
    "cm_rag_build_i16": (_i, [_fp, _i, _i, _i, _i, _fp, _sz, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _i, _i, _i, _fp, _fp]),
    "cm_node_features": (_i, [_fp, _fp, _fp, _i, _i, _i, _i, _fp, _fp, _fp, _f, _fp, _i, _fp]),
    "cm_node_features_u8": (_i, [_fp, _fp, _fp, _i, _i, _i, _i, _fp, _fp, _fp, _f, _fp, _i, _fp]),
    "cm_node_features_u8_i16": (_i, [_fp, _fp, _fp, _i, _i, _i, _i, _fp, _fp, _fp, _f, _fp, _i, _fp]),
    "cm_labels_widen_i16": (_i, [_fp, _i64, _fp, _fp]),
    "cm_edge_weights": (_i, [_fp, _fp, _fp, _fp, _i, _fp, _fp]),
    "cm_csr_workspace_bytes": (_sz, [_i, _i]),
    "cm_csr_from_edge_index": (_i, [_fp, _i64, _fp, _i, _i, _fp, _sz, _fp, _fp, _fp, _fp, _fp, _fp]),
    "cm_linear_f16": (_i, [_fp, _i, _fp, _i, _i, _i, _fp, _fp, _f, _i, _fp, _i, _fp, _fp, _fp, _fp, _fp]),
    "cm_gnn_folded_bytes": (_sz, [C.POINTER(GnnParams)]),
    "cm_gnn_fold": (_i, [C.POINTER(GnnParams), _fp, _sz, _fp]),
    "cm_gnn_workspace_bytes": (_sz, [C.POINTER(GnnParams), _i]),
    "cm_gnn_forward": (_i, [C.POINTER(GnnParams), _fp, _fp, _fp, _fp, _fp, _fp, _fp, _i, _fp, _i, _fp, _sz,
                            _fp, _fp, _fp, _fp, _fp, _fp, _fp]),
    "cm_gnn_train_workspace_bytes": (_sz, [C.POINTER(GnnParams), _i, _i]),
    "cm_train_gemm_f32": (_i, [_fp, _fp, _fp, _i, _i, _i, _i, _fp, _fp]),
    "cm_set_bn_sync": (_i, [C.c_void_p, _fp, _fp, _i]),
    "cm_gnn_train_forward": (_i, [C.POINTER(GnnParams), _fp, _fp, _fp, _fp, _fp, _i, _i, C.POINTER(GnnDropout), _fp, _sz,
                                  _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp]),
    "cm_gnn_backward": (_i, [C.POINTER(GnnParams), _fp, _fp, _fp, _fp, _fp, _i, _i, C.POINTER(GnnDropout), _fp, _sz,
                             _fp, _fp, _fp, _fp, C.POINTER(GnnGrads), _fp]),
    "cm_flash_attention_f16": (_i, [_fp, _i, _fp, _i, _fp, _i, _i, _fp, _i, _f, _fp, _i, _fp]),
    "cm_fusion_train_workspace_bytes": (_sz, [C.POINTER(FusionParams), _i, _i, _i]),
    "cm_fusion_train_forward": (_i, [C.POINTER(FusionParams), _fp, _fp, _i, _i, _fp, _i, C.POINTER(FusionDropout), _fp, _sz,
                                     _fp, _fp, _fp, _fp, _fp, _fp]),
    "cm_fusion_backward": (_i, [C.POINTER(FusionParams), _fp, _fp, _i, _i, _fp, _i, C.POINTER(FusionDropout), _fp, _sz,
                                _fp, _fp, _fp, _fp, _fp, C.POINTER(FusionGrads), _fp, _fp, _fp]),
    "cm_fusion_workspace_bytes": (_sz, [C.POINTER(FusionParams), _i, _i, _i]),
    "cm_fusion_folded_bytes": (_sz, [C.POINTER(FusionParams), _i]),
    "cm_fusion_fold": (_i, [C.POINTER(FusionParams), _fp, _i, _fp, _sz, _fp]),
    "cm_fusion_forward": (_i, [C.POINTER(FusionParams), _fp, _fp, _fp, _i, _i, _fp, _i, _i, _fp, _sz,
                               _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp]),
    "cm_canny_workspace_bytes": (_sz, [_i, _i, _i]),
    "cm_canny": (_i, [_fp, _i, _i, _i, _i, C.c_double, C.c_double, C.c_double, _fp, _sz, _fp, _fp]),
    "cm_paint_metrics": (_i, [_fp, _fp, _i, _i, _i, _fp, _fp, _fp, _i, _f, _fp, _fp, _fp, _fp]),
    "cm_paint_metrics_i16": (_i, [_fp, _fp, _i, _i, _i, _fp, _fp, _fp, _i, _f, _fp, _fp, _fp, _fp]),
    "cm_unpack_bits": (_i, [_fp, _i64, _i, _fp, _fp]),
    "cm_ingest_copy": (_i, [_fp, _fp, _i64, _i, _fp]),
    "cm_unpack_labels10": (_i, [_fp, _i64, _fp, _fp]),
    "cm_metrics_dense": (_i, [_fp, _fp, _i, _i64, _f, _fp, _fp, _fp]),
    "cm_metrics_finalize": (_i, [_fp, _fp, _i, _i64, _fp, _fp, _fp]),
    "cm_sod_workspace_bytes": (_sz, [_i, _i, _i]),
    "cm_sod_metrics": (_i, [_fp, _fp, _i, _i, _i, _i, _fp, _sz, _fp, _fp, _fp]),
    "cm_sod_metrics_static_max_bytes": (_sz, []),
    "cm_sod_metrics_static": (_i, [_fp, _fp, _i, _i, _i, _fp, _fp, _fp]),
    "cm_sod_metrics_painted": (_i, [_fp, _i, _fp, _i, _i, _i, _fp, _fp, _fp, _i, _i, _fp, _sz, _fp, _fp, _fp]),
    "cm_region_table_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "cm_region_table_build": (_i, [_fp, _i, _fp, _i, _i, _i, _i, _i, _fp, _sz, _fp]),
    "cm_region_table_metrics": (_i, [_fp, _i, _fp, _i, _i, _i, _fp, _fp, _fp, _i, _f, _fp, _sz, _fp, _fp, _fp, _fp, _fp, _fp, _fp]),
    "cm_class_confusion": (_i, [_fp, _fp, _i, _i, _fp, _fp]),
    "cm_slic_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "cm_slic": (_i, [_fp, _i, _i, _i, _i, C.c_double, C.c_double, _i, _fp, _sz, _fp, _fp]),
    "cm_slic_enforce_connectivity": (_i, [_fp, _i, _i, _i, _i, _i, _i, _fp, _fp, _fp]),
    "cm_region_targets": (_i, [_fp, _fp, _fp, _fp, _i, _i, _i, _fp, _fp, _fp, _i, _fp, _fp, _fp, _fp]),
    "cm_rg_node_loss": (_i, [_fp, _fp, _fp, _fp, _fp, _fp, _fp, _i, C.POINTER(RgLossConfig), _fp, _sz, _fp, _fp, _fp, _fp, _fp]),
    "cm_image_loss": (_i, [_fp, _fp, _fp, _fp, _fp, _fp, _fp, _i, C.POINTER(ImageLossConfig), _fp, _fp, _fp, _fp, _fp, _fp]),
    "cm_adamw_step": (_i, [_fp, _fp, _fp, _fp, _fp, _i, _i64, C.POINTER(AdamWConfig), _i, _fp, _fp, _fp]),
}

BN_SYNC_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_void_p)

_lib = None


class StreamArg(C.c_void_p):
    """A ``cudaStream_t`` argument that remembers which device it belongs to (see ``_Guarded``)."""
    device_index = None


class _Guarded:
    """The loaded library with a device guard around every call that takes a stream: the library launches on the
    CURRENT device, so a call whose stream (``stream_ptr(tensor.device)``) lives on another device makes that device
    current for the duration of the call -- what a torch op does for its own kernels.  Without it a model on cuda:1
    called while the current device is 0 would launch on a foreign stream (invalid resource handle)."""

    def __init__(self, cdll):
        self._cdll = cdll
        self._fns = {}

    def __getattr__(self, name):
        fn = self._fns.get(name)
        if fn is None:
            raw = getattr(self._cdll, name)
            takes_stream = name in SIGNATURES and len(SIGNATURES[name][1]) > 0 and SIGNATURES[name][1][-1] is _fp

            def call(*args, _raw=raw):
                st = args[-1] if args else None
                idx = getattr(st, "device_index", None)
                if idx is not None and idx != torch.cuda.current_device():
                    with torch.cuda.device(idx):
                        return _raw(*args)
                return _raw(*args)

            fn = call if takes_stream else raw
            self._fns[name] = fn
        return fn


def load() -> "_Guarded":
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `make -C camouflage_multimodal_b200/csrc` "
                "(or __graft_entry__.build()); this package has no fallback path")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)   # AttributeError if the ABI and the binding drift apart
            fn.restype = res
            fn.argtypes = args
        _lib = _Guarded(lib)
    return _lib


# Bumped by everything that writes parameters through raw device pointers (FusedAdamW.step -> cm_adamw_step): torch's
# Tensor._version does not see such writes, so every cache of folded / converted weights keys on this counter too.
_param_generation = 0


def param_generation() -> int:
    return _param_generation


def bump_param_generation(tensors=()) -> None:
    """Call after a kernel wrote ``tensors`` in place behind torch's back."""
    global _param_generation
    _param_generation += 1
    for t in tensors:
        try:
            torch.autograd.graph.increment_version(t)
        except Exception:   # the generation counter alone already invalidates this package's caches
            break


class CmError(RuntimeError):
    pass


def check(rc: int, what: str = "") -> None:
    if rc == CM_OK:
        return
    msg = load().cm_last_error().decode("utf-8", "replace")
    if rc == CM_ERR_ARG:
        raise ValueError(msg or what)
    raise CmError(f"{what}: {msg} (code {rc})")


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return None
    return C.c_void_p(t.data_ptr())


def stream_ptr(device=None):
    """The current stream of ``device`` (default: the current device) as a ``cudaStream_t`` argument."""
    st = StreamArg(torch.cuda.current_stream(device).cuda_stream)
    if device is not None:
        device = torch.device(device)
        st.device_index = device.index if device.index is not None else torch.cuda.current_device()
    return st


def require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise CmError(f"{name} lives on {t.device}; the B200 path has no CPU fallback -- move it to a CUDA device")


def f32c(t: torch.Tensor) -> torch.Tensor:
    if t.dtype == torch.float32 and t.is_contiguous():     # the usual case: one call instead of three (a training step
        return t.detach()                                  # marshals ~200 tensors)
    return t.detach().to(torch.float32).contiguous()


def timing_enable(on: bool) -> None:
    load().cm_timing_enable(1 if on else 0)


def timing_report() -> dict:
    """{kernel name: (launches, total_ms)} since the last report."""
    buf = C.create_string_buffer(1 << 16)
    load().cm_timing_report(buf, len(buf))
    out = {}
    for line in buf.value.decode().splitlines():
        name, cnt, ms = line.split()
        out[name] = (int(cnt), float(ms))
    return out
