"""CPU: the stand-in model used by the GPU drop-in tests is faithful to the reference's own `SynthesizerTrn`
(attribute names `vits_b200.dropin.config_from_model` reads, state_dict keys and shapes).  Needs the reference tree
(build container only); the stand-in's own consistency is checked everywhere."""
import json
import os
import sys
import warnings

import pytest
import torch

from oracle import fixtures as FX
from oracle import vits_oracle as O
from tests.ref_standin import SynthesizerStandIn, load_fixture
from vits_b200.dropin import config_from_model

REF = os.environ.get("VITS_REF_PATH", "/root/reference/emotional-vits")


def test_standin_matches_fixture_and_oracle(base_cfg):
    sd_wn = FX.make_state_dict(base_cfg, seed=1234, weight_norm=True)
    m = load_fixture(SynthesizerStandIn(base_cfg).eval(), sd_wn, True)
    cfg = config_from_model(m)
    assert cfg["inter_channels"] == 192 and cfg["hidden_channels"] == 256 and cfg["gin_channels"] == 1024
    assert cfg["upsample_rates"] == [8, 6, 2, 2] and cfg["upsample_kernel_sizes"] == [16, 12, 4, 4]
    assert cfg["resblock_kernel_sizes"] == [3, 7, 11] and cfg["resblock_dilation_sizes"] == [[1, 3, 5]] * 3
    inp = FX.make_p2_inputs(base_cfg, 9, seed=2)
    with torch.no_grad():
        o = m.infer_p2(inp["attn"], inp["m_p"], inp["s_p"], inp["g"], inp["noise"])
        ref = O.infer_p2(FX.make_state_dict(base_cfg, seed=1234), inp["attn"], inp["m_p"], inp["s_p"], inp["g"], inp["noise"], base_cfg)
    assert (o - ref).abs().max().item() <= 1e-5
    m.remove_weight_norm()
    assert not any(k.endswith("weight_g") for k in m.state_dict())
    with torch.no_grad():
        o2 = m.infer_p2(inp["attn"], inp["m_p"], inp["s_p"], inp["g"], inp["noise"])
    assert (o2 - ref).abs().max().item() <= 1e-5


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present (GPU box)")
def test_standin_is_faithful_to_the_reference_model(base_cfg):
    sys.path.insert(0, REF)
    warnings.filterwarnings("ignore")
    try:
        from models import SynthesizerTrn          # the reference, imported read-only
    finally:
        sys.path.remove(REF)
    cfgj = json.load(open(os.path.join(REF, "configs", "base.json")))
    real = SynthesizerTrn(cfgj["data"]["text_channels"], cfgj["data"]["filter_length"] // 2 + 1,
                          cfgj["train"]["segment_size"] // cfgj["data"]["hop_length"],
                          n_speakers=cfgj["data"]["n_speakers"], **cfgj["model"]).eval()
    stand = SynthesizerStandIn(base_cfg).eval()
    assert config_from_model(real) == config_from_model(stand)
    hot = lambda m: {k: tuple(v.shape) for k, v in m.state_dict().items() if k.startswith(("dec.", "flow."))}
    assert hot(real) == hot(stand)
    real.remove_weight_norm(); stand.remove_weight_norm()
    assert hot(real) == hot(stand)
    # the call conventions install() rebinds (models.py:219-235, 304-318, 533-534, 568-575)
    import inspect
    assert list(inspect.signature(real.flow.infer).parameters) == list(inspect.signature(stand.flow.infer).parameters)
    assert list(inspect.signature(real.flow.forward).parameters) == list(inspect.signature(stand.flow.forward).parameters)
    assert list(inspect.signature(real.dec.forward).parameters) == list(inspect.signature(stand.dec.forward).parameters)
    assert list(inspect.signature(real.infer_p2).parameters) == list(inspect.signature(stand.infer_p2).parameters)
