"""GPU: vitsb_flow_reverse_noalloc (include/vitsb200.h) - the reverse flow with its scratch in the library's static
device array - against the existing vitsb_flow_reverse, through the C ABI, on two calls in a row.

Collected last on purpose (file name): a failure here hides nothing after it.
The test is self-contained (own engine, own inputs): it passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import pytest
import torch

from oracle import fixtures as FX

pytestmark = pytest.mark.gpu

TOL_Z = 4e-3        # fp16 z tolerance of tests/test_parity_gpu.py (TOL_Z["fp16"])


def _noalloc(eng, zp, g, lens):
    B, _, T = zp.shape
    z = torch.full_like(zp, float("nan"))                       # caller-owned output
    arr = (C.c_int32 * B)(*lens) if lens is not None else None
    torch.cuda.synchronize()
    # cumulative allocation / segment counters of torch's caching allocator: they only move when something is allocated
    # (the byte totals also move when Python's garbage collector frees tensors of earlier tests during the call)
    def counters():
        st = torch.cuda.memory_stats()
        return st.get("allocation.all.allocated", 0), st.get("segment.all.allocated", 0)
    before = counters()
    rc = eng.lib.vitsb_flow_reverse_noalloc(eng._h, zp.data_ptr(), g.data_ptr(), arr, B, T, z.data_ptr(), eng._stream())
    torch.cuda.synchronize()
    assert counters() == before                                  # the call allocated nothing through torch
    return rc, z


def test_noalloc_flow_reverse_matches_existing_entry_point_twice(base_cfg, base_sd):
    from vits_b200 import FlowDecoder
    from vits_b200 import lib as L
    import numpy as np
    import os
    eng = FlowDecoder(base_cfg, base_sd, device="cuda:0", precision="fp16")
    assert eng.lib.vitsb_noalloc_scratch_bytes() == 8 << 20
    # two calls in a row with different shapes, lengths and contents: the second runs over the first call's leftovers
    cases = [([97, 40, 64], 321), ([33], 1234), ([300, 5, 224, 1], 11)]
    for lens, seed in cases:
        lat = FX.make_latents(base_cfg, lens, seed=seed)
        zp, g = lat["z_p"].cuda(), lat["g"].cuda()
        ref = eng.flow_reverse(zp, g, lens)                                   # existing entry point, caller workspace
        rc, z = _noalloc(eng, zp, g, lens)
        assert rc == 0, eng.lib.vitsb_last_error(eng._h).decode()
        assert torch.isfinite(z).all()
        assert (z - ref).abs().max().item() <= TOL_Z
        assert torch.equal(z, ref)                                            # same kernels, same order: bit-identical
        for b, Lb in enumerate(lens):
            assert z[b, :, Lb:].abs().sum().item() == 0
        rc2, z2 = _noalloc(eng, zp, g, lens)                                  # immediately again: same answer
        assert rc2 == 0 and torch.equal(z2, z)
    # unmasked form (len = NULL) against the reference-generated golden z
    gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "base_p2_T33.npz"))
    inp = FX.make_p2_inputs(base_cfg, 33, seed=1234)
    # (the .npz holds z_p Fortran-ordered; the C ABI takes contiguous [B, C, T], which FlowDecoder._prep normally ensures)
    rc, z = _noalloc(eng, torch.from_numpy(np.ascontiguousarray(gold["z_p"])).float().cuda().contiguous(),
                     inp["g"].float().cuda().contiguous(), None)
    assert rc == 0
    assert np.abs(z.cpu().numpy() - gold["z"]).max() <= TOL_Z
    # beyond the static array: rejected with a clear error, nothing launched
    big = FX.make_latents(base_cfg, [834] * 16, seed=1)
    rc, _ = _noalloc(eng, big["z_p"].cuda(), big["g"].cuda(), [834] * 16)
    assert rc == 5                                                            # VITSB_ERR_WORKSPACE
    msg = eng.lib.vitsb_last_error(eng._h).decode()
    assert "VITSB_NOALLOC_SCRATCH_BYTES" in msg and "8388608" in msg
    # a plan without the fused coupling kernel is refused, not silently run elsewhere
    e32 = FlowDecoder(base_cfg, base_sd, device="cuda:0", precision="fp32")
    lat = FX.make_latents(base_cfg, [20], seed=2)
    rc, _ = _noalloc(e32, lat["z_p"].cuda(), lat["g"].cuda(), [20])
    assert rc == 2                                                            # VITSB_ERR_UNSUPPORTED
    assert isinstance(L.STATUS[rc], str)
