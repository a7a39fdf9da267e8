import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def base_cfg():
    from vits_b200.config import HotPathConfig
    return HotPathConfig()


@pytest.fixture(scope="session")
def adapt_cfg():
    from vits_b200.config import HotPathConfig
    # sat/configs/adapt.json model block (second real shape set)
    return HotPathConfig(gin_channels=256, upsample_rates=(6, 4, 2, 2), upsample_kernel_sizes=(12, 8, 4, 4),
                         sampling_rate=8000, hop_length=96)


@pytest.fixture(scope="session")
def base_sd(base_cfg):
    from oracle import fixtures as FX
    return FX.make_state_dict(base_cfg, seed=1234, weight_norm=False)
