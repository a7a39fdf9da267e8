"""vits_b200 — B200-native (sm_100a) flow-reverse + Generator for emotional-vits inference.

Public surface:
  FlowDecoder       host-side mirror of the reference flow/decoder interface over the C ABI
  HotPathConfig     shape parameters of the path (configs/base.json `model` block)
  install(model)    drop-in rebinding behind the reference's SynthesizerTrn (vits_b200/dropin.py)
  build()           compile vits_b200/libvitsb200.so with nvcc
"""
from .build import build  # noqa: F401
from .config import HotPathConfig  # noqa: F401


def install(model, precision: str = "fp16", device=None):
    """See vits_b200.dropin.install."""
    from .dropin import install as _install
    return _install(model, precision=precision, device=device)


def config_from_model(model) -> dict:
    from .dropin import config_from_model as _c
    return _c(model)


def __getattr__(name):
    if name == "FlowDecoder":
        from .engine import FlowDecoder
        return FlowDecoder
    raise AttributeError(name)
