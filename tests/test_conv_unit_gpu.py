"""GPU: the tcgen05/TMA implicit-GEMM conv against the CUDA-core fp32 conv on
identical operand-rounded data, across every (Cin, N, k, dilation, gate)
family the model uses (swizzle 32/64/128, N tiles 32..256, multi-tile N)."""
import ctypes as C

import pytest

pytestmark = pytest.mark.gpu

CASES = [
    (0, 64, 64, 1, 1, 256, 1, 0), (0, 256, 256, 3, 1, 300, 2, 0), (0, 256, 256, 11, 5, 700, 3, 1),
    (0, 256, 512, 5, 1, 417, 2, 1), (0, 192, 512, 7, 1, 417, 1, 0), (0, 512, 2048, 3, 1, 200, 1, 0),
    (0, 256, 96, 1, 1, 417, 2, 0), (0, 128, 128, 7, 3, 1000, 2, 1), (0, 32, 64, 7, 1, 1000, 2, 0),
    (0, 32, 32, 11, 5, 2000, 2, 1), (0, 16, 32, 11, 1, 2000, 2, 0), (1, 128, 128, 7, 3, 1000, 2, 1),
    (0, 64, 128, 3, 1, 1, 1, 0), (0, 64, 128, 3, 1, 127, 2, 0), (0, 64, 128, 3, 1, 129, 2, 0),
]


@pytest.mark.parametrize("case", CASES)
def test_tc_conv_matches_f32_conv(case):
    from vits_b200 import lib as L
    lib = L.load_library()
    me, sc = C.c_double(-1), C.c_double(-1)
    rc = lib.vitsb_selftest_conv(0, *case, 4321, C.byref(me), C.byref(sc))
    assert rc == 0
    assert sc.value > 0.3
    assert me.value <= 2e-5, (case, me.value)      # same operands, fp32 accumulate: only summation order differs


# (precision, Cin, N, k, dil, L, B): 16-bit-output-only convs, i.e. the staged epilogue (swizzled shared-memory blocks +
# TMA stores), ring slots of 1-4 boxes, alternate-tile and split-column epilogue groups, odd tile counts in pair mode
STAGED = [
    (0, 64, 64, 3, 1, 1000, 3), (0, 128, 128, 3, 1, 700, 2), (0, 64, 32, 3, 1, 900, 2), (0, 64, 96, 2, 1, 500, 2),
    (0, 256, 192, 1, 1, 417, 2), (0, 128, 256, 2, 1, 300, 2), (0, 64, 128, 4, 1, 640, 1), (0, 64, 128, 5, 1, 129, 1),
    (0, 64, 64, 1, 1, 300, 1), (1, 64, 64, 3, 1, 1000, 3), (0, 512, 1024, 3, 1, 200, 2), (0, 64, 64, 3, 1, 5, 4),
]


@pytest.mark.parametrize("case", STAGED)
def test_tc_conv_staged_output_matches_f32_conv(case):
    from vits_b200 import lib as L
    lib = L.load_library()
    me, sc = C.c_double(-1), C.c_double(-1)
    rc = lib.vitsb_selftest_conv(0, *case, 2, 4321, C.byref(me), C.byref(sc))
    assert rc == 0
    assert sc.value > 0.3
    ulp = 2.0 ** -7 if case[0] == 1 else 2.0 ** -10          # bf16 / fp16 spacing at |x| in [1, 2)
    assert me.value <= ulp * max(1.0, sc.value), (case, me.value, sc.value)
