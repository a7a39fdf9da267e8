"""ctypes binding of include/vitsb200.h.  No torch types cross this boundary:
only raw pointers, sizes and the CUDA stream handle."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VITSB_LIB", os.path.join(HERE, "libvitsb200.so"))   # VITSB_LIB: A/B a differently built library
MAX_LIST = 8

PREC_FP16, PREC_BF16, PREC_FP32 = 0, 1, 2
PRECISIONS = {"fp16": PREC_FP16, "bf16": PREC_BF16, "fp32": PREC_FP32}
DT_F32, DT_F16, DT_BF16 = 0, 1, 2

STATUS = {0: "OK", 1: "BAD_ARG", 2: "UNSUPPORTED", 3: "WEIGHTS", 4: "CUDA", 5: "WORKSPACE"}


class VitsbConfig(C.Structure):
    _fields_ = [
        ("inter_channels", C.c_int32), ("hidden_channels", C.c_int32), ("gin_channels", C.c_int32),
        ("n_flows", C.c_int32), ("flow_kernel_size", C.c_int32), ("flow_n_layers", C.c_int32),
        ("flow_dilation_rate", C.c_int32 * MAX_LIST),
        ("n_resblock_kernels", C.c_int32), ("resblock_kernel_sizes", C.c_int32 * MAX_LIST),
        ("n_resblock_dilations", C.c_int32), ("resblock_dilation_sizes", (C.c_int32 * MAX_LIST) * MAX_LIST),
        ("n_upsamples", C.c_int32), ("upsample_rates", C.c_int32 * MAX_LIST),
        ("upsample_kernel_sizes", C.c_int32 * MAX_LIST), ("upsample_initial_channel", C.c_int32),
        ("precision", C.c_int32), ("enc_n_heads", C.c_int32),
    ]


class VitsbTensorDesc(C.Structure):
    _fields_ = [("name", C.c_char_p), ("data", C.c_void_p), ("dtype", C.c_int32), ("ndim", C.c_int32),
                ("shape", C.c_int64 * 4)]


# every symbol include/vitsb200.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SYMBOLS = {
    "vitsb_create": (C.c_int, [C.POINTER(VitsbConfig), C.c_int, C.POINTER(_P)]),
    "vitsb_destroy": (None, [_P]),
    "vitsb_last_error": (C.c_char_p, [_P]),
    "vitsb_set_option": (C.c_int, [_P, C.c_char_p, C.c_int]),
    "vitsb_get_option": (C.c_int, [_P, C.c_char_p, C.POINTER(C.c_int)]),
    "vitsb_load_weights": (C.c_int, [_P, C.POINTER(VitsbTensorDesc), C.c_int]),
    "vitsb_workspace_bytes": (C.c_size_t, [_P, C.c_int, C.c_int]),
    "vitsb_flow_reverse": (C.c_int, [_P, _P, _P, C.POINTER(C.c_int32), C.c_int, C.c_int, _P, _P, C.c_size_t, _P]),
    "vitsb_decode": (C.c_int, [_P, _P, _P, C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_int, _P, _P, C.c_size_t, _P]),
    "vitsb_synthesize": (C.c_int, [_P, _P, _P, C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_int, _P, _P, _P,
                                   C.c_size_t, _P]),
    "vitsb_synthesize_pcm16": (C.c_int, [_P, _P, _P, C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_int, C.c_float, _P, _P,
                                         C.c_size_t, _P]),
    "vitsb_length_regulate": (C.c_int, [_P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, _P, _P]),
    "vitsb_resample_poly": (C.c_int, [_P, _P, C.c_int, C.c_longlong, C.c_longlong, C.c_int, C.c_int, _P, C.c_int, C.c_longlong,
                                      C.c_longlong, _P, _P, C.c_float, C.c_longlong, _P]),
    "vitsb_synthesize_durations": (C.c_int, [_P, _P, _P, _P, _P, C.POINTER(C.c_int32), _P, C.c_longlong, C.c_int, C.c_int, C.c_int,
                                             C.c_float, C.c_int, _P, _P, _P, C.c_float, _P, _P, C.c_size_t, _P]),
    "vitsb_p1_workspace_bytes": (C.c_size_t, [_P, C.c_int, C.c_int]),
    "vitsb_infer_p1": (C.c_int, [_P, _P, _P, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int, C.c_int, _P, _P, _P, _P, _P,
                                 C.c_size_t, _P]),
    "vitsb_flow_reverse_noalloc": (C.c_int, [_P, _P, _P, C.POINTER(C.c_int32), C.c_int, C.c_int, _P, _P]),
    "vitsb_noalloc_scratch_bytes": (C.c_size_t, []),
    "vitsb_total_upsample": (C.c_int, [_P]),
    "vitsb_launches_per_call": (C.c_int, [_P, C.c_int, C.c_int]),
    "vitsb_flop_per_frame": (C.c_double, [_P]),
    "vitsb_profile": (C.c_int, [_P, C.c_int]),
    "vitsb_profile_read": (C.c_int, [_P, C.POINTER(C.c_int32), C.POINTER(C.c_double), C.POINTER(C.c_int64),
                                     C.POINTER(C.c_double), C.c_int]),
    "vitsb_build_info": (C.c_char_p, []),
    "vitsb_selftest_conv": (C.c_int, [C.c_int] * 9 + [C.c_uint32, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
}

_lib = None


def load_library() -> C.CDLL:
    """Load the in-tree CUDA library.  Fails loudly if it has not been built:
    there is no Python/CPU fallback for the hot path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m vits_b200.build` (nvcc, sm_100a). "
            "vits_b200 has no CPU or PyTorch fallback for the flow+decoder path.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)      # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class VitsbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"vitsb200 error {code} ({STATUS.get(code, '?')}): {msg}")
        self.code = code


def make_config(cfg, precision: str) -> VitsbConfig:
    """cfg: any object with the hot-path attribute names of configs/base.json
    (e.g. the reference HParams `hps.model`, a dict, or a dataclass)."""
    def get(name, default=None):
        if isinstance(cfg, dict):
            return cfg.get(name, default)
        return getattr(cfg, name, default)

    c = VitsbConfig()
    c.inter_channels = get("inter_channels")
    c.hidden_channels = get("hidden_channels")
    c.gin_channels = get("gin_channels")
    c.n_flows = get("n_flows", 4)
    c.flow_kernel_size = get("flow_kernel_size", 5)      # models.py:462 hard-codes 5
    c.flow_n_layers = get("flow_n_layers", 4)            # models.py:462 hard-codes 4
    dr = list(get("dilation_rate", [1] * c.n_flows))
    for i, v in enumerate(dr[:MAX_LIST]):
        c.flow_dilation_rate[i] = v
    ks = list(get("resblock_kernel_sizes"))
    ds = [list(d) for d in get("resblock_dilation_sizes")]
    if str(get("resblock", "2")) != "2":
        raise VitsbError(2, "only resblock '2' is supported (ResBlock1 is shape-inconsistent in the reference, "
                            "modules.py:185-220)")
    if len(ks) > MAX_LIST or len(ds) != len(ks) or any(len(d) != len(ds[0]) or len(d) > MAX_LIST for d in ds):
        raise VitsbError(2, "resblock kernel/dilation lists")
    c.n_resblock_kernels = len(ks)
    c.n_resblock_dilations = len(ds[0])
    for j, k in enumerate(ks):
        c.resblock_kernel_sizes[j] = k
        for d, v in enumerate(ds[j]):
            c.resblock_dilation_sizes[j][d] = v
    ur, uk = list(get("upsample_rates")), list(get("upsample_kernel_sizes"))
    if len(ur) != len(uk) or len(ur) > MAX_LIST:
        raise VitsbError(2, "upsample lists")
    c.n_upsamples = len(ur)
    for i, (u, k) in enumerate(zip(ur, uk)):
        c.upsample_rates[i] = u
        c.upsample_kernel_sizes[i] = k
    c.upsample_initial_channel = get("upsample_initial_channel")
    c.precision = PRECISIONS[precision]
    c.enc_n_heads = int(get("n_heads", 2) or 2)
    return c
