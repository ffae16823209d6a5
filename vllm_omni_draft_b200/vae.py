"""VAE decode and encode for the image pipeline — host layer.

What the reference reaches through ``self.vae.decode(latents)`` (src/qwen_image/runner_adapter.py:149-160,
src/qwen_image_gen/model.py:138-141) and ``pipe(...).images`` (test_performance.py:277): diffusers'
``AutoencoderKLQwenImage`` decoder.  The image pipeline decodes exactly ONE latent frame, and for one frame
every causal 3-D convolution of that decoder degenerates exactly:

* time padding is 2·p zeros in front and none behind, so a (3,k,k) kernel only ever multiplies its LAST
  temporal slice with real data  ->  a plain 2-D k×k convolution with ``weight[:, :, -1]``;
* the temporal up-sampling conv of 'upsample3d' is skipped for the first chunk of a clip  ->  both resample
  modes are nearest 2× + 3×3 conv.

So this module stores 2-D weights (a third of the checkpoint's conv bytes), keeps activations channels-last
in bf16, and runs as ordinary library convolutions: it is 0.05 % of an image's FLOPs (≈4 TF against 7 830 TF
for 50 true-CFG steps at 1024²) and sits outside the MMDiT block, whose kernels are the hand-written part
of this package.  tests/ check it against the 5-D causal restatement in oracle/vae_oracle.py.

The ENCODER (``B200QwenImageVAE.encode``; the edit pipeline's ``vae.encode(image)`` upstream, reached by the reference
through ``pipe(image=<PIL images>, ...)``, src/performance_enhancement/test_performance.py:244-251) reduces the same
way for one frame: every causal conv keeps its last temporal slice, and the stride-2 temporal conv of
'downsample3d' only runs from a clip's second chunk on, so both resample modes are ZeroPad2d((0,1,0,1)) + a 3x3
stride-2 conv.
"""
from __future__ import annotations

import glob
import os
from dataclasses import dataclass
from types import SimpleNamespace
from typing import Dict, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

LATENTS_MEAN = (-0.7571, -0.7089, -0.9113, 0.1075, -0.1745, 0.9653, -0.1517, 1.5508,
                0.4134, -0.0715, 0.5517, -0.3632, -0.1922, -0.9497, 0.2503, -0.2921)
LATENTS_STD = (2.8184, 1.4541, 2.3275, 2.6558, 1.2196, 1.7708, 2.6052, 2.0743,
               3.2687, 2.1526, 2.8652, 1.5579, 1.6382, 1.1253, 2.8251, 1.9160)


@dataclass
class DecoderOutput:
    sample: torch.Tensor


class _ChannelNorm(nn.Module):
    """x / max(||x||_2 over channels, 1e-12) * sqrt(C) * gamma  (QwenImageRMS_norm, channel first)."""

    def __init__(self, dim: int):
        super().__init__()
        self.gamma = nn.Parameter(torch.ones(dim, 1, 1), requires_grad=False)
        self.scale = dim ** 0.5

    def forward(self, x):
        return F.normalize(x, dim=1) * self.scale * self.gamma


class _ResBlock(nn.Module):
    def __init__(self, cin: int, cout: int):
        super().__init__()
        self.norm1, self.conv1 = _ChannelNorm(cin), nn.Conv2d(cin, cout, 3, padding=1)
        self.norm2, self.conv2 = _ChannelNorm(cout), nn.Conv2d(cout, cout, 3, padding=1)
        self.conv_shortcut = nn.Conv2d(cin, cout, 1) if cin != cout else None

    def forward(self, x):
        h = x if self.conv_shortcut is None else self.conv_shortcut(x)
        x = self.conv1(F.silu(self.norm1(x)))
        x = self.conv2(F.silu(self.norm2(x)))
        return x + h


class _SpatialAttention(nn.Module):
    """single-head attention over the H·W positions (mid block)."""

    def __init__(self, dim: int):
        super().__init__()
        self.norm = _ChannelNorm(dim)
        self.to_qkv = nn.Conv2d(dim, 3 * dim, 1)
        self.proj = nn.Conv2d(dim, dim, 1)

    def forward(self, x):
        b, c, h, w = x.shape
        qkv = self.to_qkv(self.norm(x)).reshape(b, 1, 3 * c, h * w).transpose(2, 3)
        q, k, v = qkv.chunk(3, dim=-1)
        o = F.scaled_dot_product_attention(q, k, v)
        o = o.squeeze(1).transpose(1, 2).reshape(b, c, h, w)
        return x + self.proj(o)


class _Up2x(nn.Module):
    def __init__(self, dim: int):
        super().__init__()
        self.conv = nn.Conv2d(dim, dim // 2, 3, padding=1)

    def forward(self, x):
        # nearest-exact with an integer factor 2 is plain pixel replication
        x = x.repeat_interleave(2, dim=2).repeat_interleave(2, dim=3)
        return self.conv(x)


class _Down2x(nn.Module):
    def __init__(self, dim: int):
        super().__init__()
        self.conv = nn.Conv2d(dim, dim, 3, stride=2)

    def forward(self, x):
        return self.conv(F.pad(x, (0, 1, 0, 1)))


class DiagonalGaussian:
    """The slice of diffusers' DiagonalGaussianDistribution the pipelines use (mode / mean / sample)."""

    def __init__(self, moments: torch.Tensor):
        self.mean, self.logvar = moments.chunk(2, dim=1)
        self.logvar = self.logvar.clamp(-30.0, 20.0)

    def mode(self) -> torch.Tensor:
        return self.mean

    def sample(self, generator: Optional[torch.Generator] = None) -> torch.Tensor:
        noise = torch.randn(self.mean.shape, generator=generator, device=self.mean.device, dtype=self.mean.dtype)
        return self.mean + torch.exp(0.5 * self.logvar) * noise


@dataclass
class EncoderOutput:
    latent_dist: DiagonalGaussian


class B200QwenImageVAEDecoder(nn.Module):
    def __init__(self, base_dim: int = 96, z_dim: int = 16, dim_mult: Tuple[int, ...] = (1, 2, 4, 4),
                 num_res_blocks: int = 2, latents_mean=LATENTS_MEAN, latents_std=LATENTS_STD,
                 device="cuda", dtype=torch.bfloat16):
        super().__init__()
        self.config = SimpleNamespace(base_dim=base_dim, z_dim=z_dim, dim_mult=tuple(dim_mult),
                                      num_res_blocks=num_res_blocks, latents_mean=list(latents_mean),
                                      latents_std=list(latents_std))
        dims = [base_dim * u for u in (dim_mult[-1],) + tuple(reversed(dim_mult))]
        self.post_quant_conv = nn.Conv2d(z_dim, z_dim, 1)
        self.conv_in = nn.Conv2d(z_dim, dims[0], 3, padding=1)
        self.mid = nn.ModuleList([_ResBlock(dims[0], dims[0]), _SpatialAttention(dims[0]),
                                  _ResBlock(dims[0], dims[0])])
        self.up_res = nn.ModuleList()
        self.up_sample = nn.ModuleList()
        for i, (cin, cout) in enumerate(zip(dims[:-1], dims[1:])):
            if i > 0:
                cin //= 2
            blocks = nn.ModuleList()
            for _ in range(num_res_blocks + 1):
                blocks.append(_ResBlock(cin, cout))
                cin = cout
            self.up_res.append(blocks)
            self.up_sample.append(_Up2x(cout) if i != len(dim_mult) - 1 else nn.Identity())
        self.norm_out = _ChannelNorm(dims[-1])
        self.conv_out = nn.Conv2d(dims[-1], 3, 3, padding=1)
        self.requires_grad_(False)
        self.to(device=device, dtype=dtype, memory_format=torch.channels_last)

    @property
    def dtype(self):
        return self.conv_in.weight.dtype

    @property
    def device(self):
        return self.conv_in.weight.device

    # ------------------------------------------------------------------ weights
    @torch.no_grad()
    def load_diffusers_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        """AutoencoderKLQwenImage names -> 2-D weights.  3-D kernels keep only their last temporal slice
        (the only one that meets real data for a single frame); encoder / time_conv entries are unused."""
        self._load_decoder(sd)

    @staticmethod
    def _loaders(sd):
        def conv(m: nn.Conv2d, name: str):
            w = sd[name + ".weight"]
            if w.dim() == 5:
                w = w[:, :, -1]
            m.weight.copy_(w.to(m.weight.dtype))
            m.bias.copy_(sd[name + ".bias"].to(m.bias.dtype))

        def gamma(m: _ChannelNorm, name: str):
            m.gamma.copy_(sd[name + ".gamma"].reshape(-1, 1, 1).to(m.gamma.dtype))

        def res(m: _ResBlock, p: str):
            gamma(m.norm1, p + "norm1"); conv(m.conv1, p + "conv1")
            gamma(m.norm2, p + "norm2"); conv(m.conv2, p + "conv2")
            if m.conv_shortcut is not None:
                conv(m.conv_shortcut, p + "conv_shortcut")

        return conv, gamma, res

    @torch.no_grad()
    def _load_decoder(self, sd: Dict[str, torch.Tensor]) -> None:
        conv, gamma, res = self._loaders(sd)
        conv(self.post_quant_conv, "post_quant_conv")
        conv(self.conv_in, "decoder.conv_in")
        res(self.mid[0], "decoder.mid_block.resnets.0.")
        a = "decoder.mid_block.attentions.0."
        gamma(self.mid[1].norm, a + "norm"); conv(self.mid[1].to_qkv, a + "to_qkv"); conv(self.mid[1].proj, a + "proj")
        res(self.mid[2], "decoder.mid_block.resnets.1.")
        for i, blocks in enumerate(self.up_res):
            for j, m in enumerate(blocks):
                res(m, f"decoder.up_blocks.{i}.resnets.{j}.")
            if isinstance(self.up_sample[i], _Up2x):
                conv(self.up_sample[i].conv, f"decoder.up_blocks.{i}.upsamplers.0.resample.1")
        gamma(self.norm_out, "decoder.norm_out")
        conv(self.conv_out, "decoder.conv_out")

    @classmethod
    def from_pretrained(cls, model_id, subfolder: Optional[str] = None, torch_dtype=torch.bfloat16,
                        device_map=None, device="cuda"):
        """Local directory with diffusers-named ``*.safetensors`` (reference: model.py:59-65); no hub access."""
        path = os.path.join(model_id, subfolder) if subfolder else model_id
        files = sorted(glob.glob(os.path.join(path, "*.safetensors")))
        if not files:
            raise RuntimeError(f"Failed to load QwenImage components: no safetensors under {path!r}")
        from safetensors.torch import load_file
        sd = {}
        for f in files:
            sd.update(load_file(f))
        m = cls(device=device, dtype=torch_dtype)
        m.load_diffusers_state_dict(sd)
        return m

    def normalize(self, latents: torch.Tensor) -> torch.Tensor:
        """The edit pipeline's ``(latents - mean) / std`` (per latent channel) after encode."""
        shape = [1, self.config.z_dim] + [1] * (latents.dim() - 2)
        mean = torch.tensor(self.config.latents_mean, device=latents.device, dtype=latents.dtype).view(shape)
        std = torch.tensor(self.config.latents_std, device=latents.device, dtype=latents.dtype).view(shape)
        return (latents - mean) / std

    # ------------------------------------------------------------------ decode
    def denormalize(self, latents: torch.Tensor) -> torch.Tensor:
        """The pipeline's ``latents / (1/std) + mean`` (per latent channel) ahead of decode."""
        shape = [1, self.config.z_dim] + [1] * (latents.dim() - 2)
        mean = torch.tensor(self.config.latents_mean, device=latents.device, dtype=latents.dtype).view(shape)
        inv_std = 1.0 / torch.tensor(self.config.latents_std, device=latents.device, dtype=latents.dtype).view(shape)
        return latents / inv_std + mean

    @torch.no_grad()
    def decode(self, z: torch.Tensor, return_dict: bool = True):
        """z: [B, z_dim, 1, H, W] (or [B, z_dim, H, W]) -> sample [B, 3, 1, 8H, 8W] clamped to [-1, 1]."""
        five_d = z.dim() == 5
        if five_d:
            if z.shape[2] != 1:
                raise ValueError("the image pipeline decodes exactly one latent frame")
            z = z[:, :, 0]
        if z.dim() != 4 or z.shape[1] != self.config.z_dim:
            raise ValueError(f"latents must be [B, {self.config.z_dim}, (1,) H, W]")
        x = z.to(self.device, self.dtype).contiguous(memory_format=torch.channels_last)
        x = self.conv_in(self.post_quant_conv(x))
        for m in self.mid:
            x = m(x)
        for blocks, up in zip(self.up_res, self.up_sample):
            for m in blocks:
                x = m(x)
            x = up(x)
        x = self.conv_out(F.silu(self.norm_out(x)))
        x = torch.clamp(x, min=-1.0, max=1.0).unsqueeze(2)
        if not five_d:
            x = x[:, :, 0]
        if not return_dict:
            return (x,)
        return DecoderOutput(sample=x)


class B200QwenImageVAE(B200QwenImageVAEDecoder):
    """Decoder + one-frame encoder (``AutoencoderKLQwenImage.encode`` / ``.decode`` for images)."""

    def __init__(self, base_dim: int = 96, z_dim: int = 16, dim_mult: Tuple[int, ...] = (1, 2, 4, 4),
                 num_res_blocks: int = 2, latents_mean=LATENTS_MEAN, latents_std=LATENTS_STD,
                 device="cuda", dtype=torch.bfloat16):
        super().__init__(base_dim, z_dim, dim_mult, num_res_blocks, latents_mean, latents_std, device, dtype)
        dims = [base_dim * u for u in (1,) + tuple(dim_mult)]
        self.enc_conv_in = nn.Conv2d(3, dims[0], 3, padding=1)
        self.enc_down = nn.ModuleList()         # flat, like upstream's down_blocks: residual blocks and resamples
        for i, (cin, cout) in enumerate(zip(dims[:-1], dims[1:])):
            for _ in range(num_res_blocks):
                self.enc_down.append(_ResBlock(cin, cout))
                cin = cout
            if i != len(dim_mult) - 1:
                self.enc_down.append(_Down2x(cout))
        top = dims[-1]
        self.enc_mid = nn.ModuleList([_ResBlock(top, top), _SpatialAttention(top), _ResBlock(top, top)])
        self.enc_norm_out = _ChannelNorm(top)
        self.enc_conv_out = nn.Conv2d(top, 2 * z_dim, 3, padding=1)
        self.quant_conv = nn.Conv2d(2 * z_dim, 2 * z_dim, 1)
        self.requires_grad_(False)
        self.to(device=device, dtype=dtype, memory_format=torch.channels_last)

    @torch.no_grad()
    def load_diffusers_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        self._load_decoder(sd)
        conv, gamma, res = self._loaders(sd)
        conv(self.enc_conv_in, "encoder.conv_in")
        for k, m in enumerate(self.enc_down):
            if isinstance(m, _ResBlock):
                res(m, f"encoder.down_blocks.{k}.")
            else:
                conv(m.conv, f"encoder.down_blocks.{k}.resample.1")     # time_conv: never met by a single frame
        res(self.enc_mid[0], "encoder.mid_block.resnets.0.")
        a = "encoder.mid_block.attentions.0."
        gamma(self.enc_mid[1].norm, a + "norm"); conv(self.enc_mid[1].to_qkv, a + "to_qkv")
        conv(self.enc_mid[1].proj, a + "proj")
        res(self.enc_mid[2], "encoder.mid_block.resnets.1.")
        gamma(self.enc_norm_out, "encoder.norm_out")
        conv(self.enc_conv_out, "encoder.conv_out")
        conv(self.quant_conv, "quant_conv")

    @torch.no_grad()
    def encode(self, x: torch.Tensor, return_dict: bool = True):
        """x: [B, 3, 1, H, W] (or [B, 3, H, W]) in [-1, 1], H and W multiples of 8 -> posterior over
        [B, z_dim, 1, H/8, W/8] (``.latent_dist.mode()`` is what the edit pipeline takes)."""
        five_d = x.dim() == 5
        if five_d:
            if x.shape[2] != 1:
                raise ValueError("the image pipeline encodes exactly one frame")
            x = x[:, :, 0]
        if x.dim() != 4 or x.shape[1] != 3 or x.shape[2] % 8 or x.shape[3] % 8:
            raise ValueError("images must be [B, 3, (1,) H, W] with H and W multiples of 8")
        x = x.to(self.device, self.dtype).contiguous(memory_format=torch.channels_last)
        x = self.enc_conv_in(x)
        for m in self.enc_down:
            x = m(x)
        for m in self.enc_mid:
            x = m(x)
        x = self.quant_conv(self.enc_conv_out(F.silu(self.enc_norm_out(x))))
        if five_d:
            x = x.unsqueeze(2)
        dist = DiagonalGaussian(x)
        if not return_dict:
            return (dist,)
        return EncoderOutput(latent_dist=dist)
