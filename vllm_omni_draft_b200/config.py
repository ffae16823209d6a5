"""Configuration of the hot path.  The reference's own config (src/qwen_image_gen/config.py:10-66)
is an env-var dataclass for the vLLM wrapper; only the fields the denoising path needs are kept."""
from __future__ import annotations

import os
from dataclasses import dataclass, fields
from typing import Tuple


@dataclass(frozen=True)
class MMDiTConfig:
    """QwenImageTransformer2DModel hyper-parameters."""
    num_layers: int = 60
    attention_head_dim: int = 128
    num_attention_heads: int = 24
    joint_attention_dim: int = 3584
    in_channels: int = 64
    out_channels: int = 16
    patch_size: int = 2
    axes_dims_rope: Tuple[int, int, int] = (16, 56, 56)
    rope_theta: float = 10000.0
    guidance_embeds: bool = False

    @property
    def inner_dim(self) -> int:
        return self.num_attention_heads * self.attention_head_dim

    @property
    def proj_out_dim(self) -> int:
        return self.patch_size * self.patch_size * self.out_channels

    @staticmethod
    def tiny() -> "MMDiTConfig":
        """BASELINE.json configs[0]."""
        return MMDiTConfig(num_layers=2, attention_head_dim=64, num_attention_heads=4,
                           joint_attention_dim=128, axes_dims_rope=(8, 28, 28))

    @staticmethod
    def qwen_image_20b() -> "MMDiTConfig":
        return MMDiTConfig()


@dataclass
class RunConfig:
    """Per-run knobs; env overrides follow the reference's QWEN_* names (config.py:44-57)."""
    height: int = 1024
    width: int = 1024
    num_inference_steps: int = 50
    true_cfg_scale: float = 4.0
    txt_len: int = 512
    batch: int = 1
    seed: int = 42
    cfg_parallel: int = 1      # 1 or 2 ranks for cond/uncond
    sp_degree: int = 1         # Ulysses degree

    @staticmethod
    def from_env(**kwargs) -> "RunConfig":
        env = {"height": "QWEN_HEIGHT", "width": "QWEN_WIDTH", "num_inference_steps": "QWEN_NUM_STEPS",
               "true_cfg_scale": "QWEN_GUIDANCE_SCALE", "batch": "QWEN_MAX_BATCH_SIZE"}
        cfg = RunConfig()
        for f, var in env.items():
            if var in os.environ:
                setattr(cfg, f, type(getattr(cfg, f))(os.environ[var]))
        names = {f.name for f in fields(RunConfig)}
        for k, v in kwargs.items():
            if k not in names:
                raise ValueError(f"Unknown config parameter: {k}")   # config.py:60-64
            setattr(cfg, k, v)
        return cfg
