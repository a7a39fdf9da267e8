"""Shape parameters of the flow + decoder hot path (the `model` block of the reference's configs/base.json).

This dataclass is the product's public config type: `FlowDecoder(cfg, state_dict)` accepts it (or the reference's
`hps.model`, or a dict with the same names); the test oracle imports it from here.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Sequence


@dataclass
class HotPathConfig:
    """Shape parameters of the hot path (configs/base.json `model` block)."""
    inter_channels: int = 192
    hidden_channels: int = 256
    gin_channels: int = 1024
    n_flows: int = 4
    flow_kernel_size: int = 5          # models.py:462 (hard-coded 5)
    flow_n_layers: int = 4             # models.py:462 (hard-coded 4)
    dilation_rate: Sequence[int] = (1, 1, 1, 1)
    resblock_kernel_sizes: Sequence[int] = (3, 7, 11)
    resblock_dilation_sizes: Sequence[Sequence[int]] = ((1, 3, 5), (1, 3, 5), (1, 3, 5))
    upsample_rates: Sequence[int] = (8, 6, 2, 2)
    upsample_initial_channel: int = 512
    upsample_kernel_sizes: Sequence[int] = (16, 12, 4, 4)
    sampling_rate: int = 16000
    hop_length: int = 192
    # infer_p1 side (text encoder + duration predictor; models.py:421-448, 462-465)
    text_channels: int = 256
    filter_channels: int = 512
    n_heads: int = 2
    n_layers: int = 6
    kernel_size: int = 5
    hidden_size_d: int = 256
    kernel_size_d: int = 5
    n_speakers: int = 2048
    emo_channels: int = 1024           # models.py:147 hard-codes nn.Linear(1024, hidden_channels)

    @staticmethod
    def from_json(cfg: dict) -> "HotPathConfig":
        m, d = cfg["model"], cfg.get("data", {})
        return HotPathConfig(
            inter_channels=m["inter_channels"],
            hidden_channels=m["hidden_channels"],
            gin_channels=m["gin_channels"],
            n_flows=m.get("n_flows", 4),
            dilation_rate=tuple(m.get("dilation_rate", [1, 1, 1, 1])),
            resblock_kernel_sizes=tuple(m["resblock_kernel_sizes"]),
            resblock_dilation_sizes=tuple(tuple(x) for x in m["resblock_dilation_sizes"]),
            upsample_rates=tuple(m["upsample_rates"]),
            upsample_initial_channel=m["upsample_initial_channel"],
            upsample_kernel_sizes=tuple(m["upsample_kernel_sizes"]),
            sampling_rate=d.get("sampling_rate", 16000),
            hop_length=d.get("hop_length", 192),
            text_channels=d.get("text_channels", 256),
            filter_channels=m.get("filter_channels", 512),
            n_heads=m.get("n_heads", 2),
            n_layers=m.get("n_layers", 6),
            kernel_size=m.get("kernel_size", 5),
            hidden_size_d=m.get("hidden_size_d", 256),
            kernel_size_d=m.get("kernel_size_d", 5),
            n_speakers=d.get("n_speakers", 2048),
        )

    @property
    def total_upsample(self) -> int:
        r = 1
        for u in self.upsample_rates:
            r *= u
        return r
