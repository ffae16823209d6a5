"""MSRoPE tables for the fused QK-norm+RoPE kernel (host layer).

Integer position ids come from the C ABI (`qi_rope_position_ids`, bit-exact restatement of
QwenEmbedRope's index arithmetic); the rotation angles are pos * theta^(-2j/a) in fp32 and the
(cos, sin) pairs are laid out [S_joint, Dh/2, 2] with text rows first, matching the joint
[txt; img] token order of the attention."""
from __future__ import annotations

from functools import lru_cache
from typing import Sequence, Tuple

import torch

from . import ops


def normalize_img_shapes(img_shapes):
    """Same acceptance rules as the upstream pos_embed: list -> [0]; bare tuple -> [tuple]."""
    if isinstance(img_shapes, list):
        img_shapes = img_shapes[0]
    if not isinstance(img_shapes, list):
        img_shapes = [img_shapes]
    return tuple(tuple(int(v) for v in s) for s in img_shapes)


@lru_cache(maxsize=16)
def _joint_cos_sin_cached(axes: Tuple[int, ...], theta: float, shapes: Tuple[Tuple[int, int, int], ...],
                          txt_len: int) -> torch.Tensor:
    img_pos, txt_pos = ops.rope_position_ids(shapes, txt_len)
    pos = torch.cat([txt_pos, img_pos], dim=0).to(torch.int64)          # text first
    parts = []
    for a, dim in enumerate(axes):
        omega = 1.0 / torch.pow(theta, torch.arange(0, dim, 2).to(torch.float32).div(dim))
        ang = torch.outer(pos[:, a], omega)                              # int64 x fp32 -> fp32
        parts.append(ang)
    ang = torch.cat(parts, dim=1)
    # torch.polar is what upstream uses; going through it keeps the table bit-identical
    return torch.view_as_real(torch.polar(torch.ones_like(ang), ang)).contiguous()


def joint_cos_sin(axes: Sequence[int], theta: float, img_shapes, txt_len: int) -> torch.Tensor:
    shapes = tuple(tuple(int(v) for v in s) for s in img_shapes)
    return _joint_cos_sin_cached(tuple(int(a) for a in axes), float(theta), shapes, int(txt_len))
