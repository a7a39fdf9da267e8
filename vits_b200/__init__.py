"""vits_b200 — B200-native (sm_100a) flow-reverse + Generator for emotional-vits inference.

Public surface:
  FlowDecoder       host-side mirror of the reference flow/decoder interface over the C ABI
  install(model)    drop-in rebinding behind the reference's SynthesizerTrn
  build()           compile vits_b200/libvitsb200.so with nvcc
"""
from .build import build  # noqa: F401


def __getattr__(name):
    if name in ("FlowDecoder",):
        from .engine import FlowDecoder
        return FlowDecoder
    if name in ("install", "config_from_model"):
        from . import install as _i
        return getattr(_i, name)
    raise AttributeError(name)
