"""CPU: the C-ABI library loads, exports every symbol include/vitsb200.h declares,
validates configs without a device, and refuses to compute without one (no fallback)."""
import ctypes as C
import os
import re

import pytest
import torch

import __graft_entry__ as entry
from vits_b200 import lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    entry.build()
    return L.load_library()


def test_header_symbols_all_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "vitsb200.h")).read()
    declared = set(re.findall(r"\b(vitsb_[a-z_0-9]+)\s*\(", hdr))
    declared -= {"vitsb_status", "vitsb_precision"}
    assert declared == set(L.SYMBOLS), declared ^ set(L.SYMBOLS)
    for name in declared:
        assert getattr(lib, name) is not None


def test_build_info_is_blackwell_native(lib):
    info = lib.vitsb_build_info().decode()
    assert "sm_100a" in info and "tcgen05" in info and "cpu_fallback=none" in info


def test_sass_has_tcgen05_and_tma():
    import shutil
    import subprocess
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not available")
    sass = subprocess.run(["cuobjdump", "-sass", L.LIB_PATH], capture_output=True, text=True).stdout
    assert "UTCHMMA" in sass or "UTCMMA" in sass      # tcgen05.mma
    assert "UTMALDG" in sass                           # TMA tensor load
    assert "LDTM" in sass                              # tcgen05.ld
    import re
    assert "UTCHMMA.2CTA" in sass                      # CTA-pair (cta_group::2) MMAs
    assert not re.search(r"(?<![A-Z])HMMA\.", sass)    # no legacy mma.sync path (UTCHMMA.2CTA contains the substring)


def test_config_validation_without_device(lib, base_cfg):
    h = C.c_void_p()
    c = L.make_config(base_cfg, "fp16")
    c.upsample_initial_channel = 100
    assert lib.vitsb_create(C.byref(c), 0, C.byref(h)) == 2          # UNSUPPORTED
    assert b"upsample_initial_channel" in lib.vitsb_last_error(None)
    c = L.make_config(base_cfg, "fp16")
    c.precision = 9
    assert lib.vitsb_create(C.byref(c), 0, C.byref(h)) == 2
    with pytest.raises(L.VitsbError):
        L.make_config({**base_cfg.__dict__, "resblock": "1"}, "fp16")   # ResBlock1: broken in the reference
    c = L.make_config({**base_cfg.__dict__, "n_flows": 3, "dilation_rate": [1, 1, 1]}, "fp16")
    assert lib.vitsb_create(C.byref(c), 0, C.byref(h)) == 2          # odd n_flows: the folded Flips would not cancel
    assert b"n_flows" in lib.vitsb_last_error(None)


def test_no_environment_switches_in_the_library():
    """Launch-plan choices are vitsb_set_option keys; getenv survives only inside the debug-hook build."""
    import re
    csrc = os.path.join(ROOT, "vits_b200", "csrc")
    for f in os.listdir(csrc):
        src = open(os.path.join(csrc, f)).read()
        src = re.sub(r"#ifdef VITSB_RB_DEBUG.*?#endif", "", src, flags=re.S)
        assert "getenv" not in src, f


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback(lib, base_cfg, base_sd):
    h = C.c_void_p()
    c = L.make_config(base_cfg, "fp32")
    assert lib.vitsb_create(C.byref(c), 0, C.byref(h)) == 4          # ERR_CUDA: no device, no fallback
    from vits_b200 import FlowDecoder
    with pytest.raises(RuntimeError):
        FlowDecoder(base_cfg, base_sd, device="cuda:0")


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "vits_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
