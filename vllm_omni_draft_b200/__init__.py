"""B200-native Qwen-Image denoising hot path (drop-in behind the reference's transformer forward and
pipeline call signatures)."""
from .config import MMDiTConfig, RunConfig  # noqa: F401

__all__ = ["MMDiTConfig", "RunConfig"]
