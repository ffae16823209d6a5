"""ORACLE — test infrastructure only.  PARITY UNPINNED.

Plain-PyTorch restatement of the *decoder half* of diffusers' ``AutoencoderKLQwenImage`` (the Wan-2.1
style causal-3D-conv VAE) plus the pipeline's latent de-normalisation, i.e. what the reference reaches
through ``self.vae.decode(latents)`` (reference src/qwen_image/runner_adapter.py:149-160,
src/qwen_image_gen/model.py:138-141) and through ``pipe(...).images`` (test_performance.py:277).  The class
itself is in the un-vendored ``diffusers`` package, so — like qwen_image_oracle.py — this is a restatement
of the published algorithm with no golden vectors to pin it: **parity unpinned** (SURVEY.md §8 row f1).

Written the way upstream executes it: 5-D tensors ``[B, C, T, H, W]``, every convolution a *causal* 3-D
convolution (zero padding 2·p in front of the time axis, none behind), functional style over a state dict
with diffusers parameter names.  The product's decoder (vllm_omni_draft_b200/vae.py) is specialised to one
frame and 2-D convolutions; tests check the two agree, which is what validates that reduction.

Only ``tests/`` may import this file.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Sequence, Tuple

import torch
import torch.nn.functional as F

# AutoencoderKLQwenImage config defaults (same statistics as Wan 2.1)
LATENTS_MEAN = (-0.7571, -0.7089, -0.9113, 0.1075, -0.1745, 0.9653, -0.1517, 1.5508,
                0.4134, -0.0715, 0.5517, -0.3632, -0.1922, -0.9497, 0.2503, -0.2921)
LATENTS_STD = (2.8184, 1.4541, 2.3275, 2.6558, 1.2196, 1.7708, 2.6052, 2.0743,
               3.2687, 2.1526, 2.8652, 1.5579, 1.6382, 1.1253, 2.8251, 1.9160)


@dataclass(frozen=True)
class VAEOracleConfig:
    base_dim: int = 96
    z_dim: int = 16
    dim_mult: Tuple[int, ...] = (1, 2, 4, 4)
    num_res_blocks: int = 2
    temperal_downsample: Tuple[bool, ...] = (False, True, True)
    latents_mean: Tuple[float, ...] = LATENTS_MEAN
    latents_std: Tuple[float, ...] = LATENTS_STD

    def decoder_dims(self):
        """[(in_dim, out_dim, upsample_mode or None)] per up block (QwenImageDecoder3d.__init__)."""
        dims = [self.base_dim * u for u in (self.dim_mult[-1],) + tuple(reversed(self.dim_mult))]
        t_up = tuple(reversed(self.temperal_downsample))
        out = []
        for i, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
            if i > 0:
                a = a // 2
            mode = None
            if i != len(self.dim_mult) - 1:
                mode = "upsample3d" if t_up[i] else "upsample2d"
            out.append((a, b, mode))
        return dims[0], out


def encoder_dims(cfg: VAEOracleConfig):
    """[(in_dim, out_dim, downsample_mode or None)] per encoder stage (QwenImageEncoder3d.__init__:
    dims = base_dim * [1] + dim_mult; 'downsample3d' where temperal_downsample says so)."""
    dims = [cfg.base_dim * u for u in (1,) + tuple(cfg.dim_mult)]
    out = []
    for i, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
        mode = None
        if i != len(cfg.dim_mult) - 1:
            mode = "downsample3d" if cfg.temperal_downsample[i] else "downsample2d"
        out.append((a, b, mode))
    return out


def tiny_vae_config() -> VAEOracleConfig:
    return VAEOracleConfig(base_dim=8, dim_mult=(1, 2, 4, 4), num_res_blocks=1)


def init_vae_state_dict(cfg: VAEOracleConfig, seed: int = 0, weight_std: float = 0.05,
                        bias_std: float = 0.02, gamma_std: float = 0.1,
                        with_encoder: bool = True) -> Dict[str, torch.Tensor]:
    """Random decoder (+ encoder) weights under diffusers parameter names (fp32).  The encoder entries are drawn
    AFTER the decoder's, so the decoder weights of a seed do not depend on `with_encoder` (golden fixtures)."""
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, torch.Tensor] = {}

    def conv3(name, o, i, k):
        kt, kh, kw = k
        sd[name + ".weight"] = torch.randn(o, i, kt, kh, kw, generator=g) * weight_std
        sd[name + ".bias"] = torch.randn(o, generator=g) * bias_std

    def conv2(name, o, i, k):
        sd[name + ".weight"] = torch.randn(o, i, k, k, generator=g) * weight_std
        sd[name + ".bias"] = torch.randn(o, generator=g) * bias_std

    def gamma(name, dim, images):
        shape = (dim, 1, 1) if images else (dim, 1, 1, 1)
        sd[name + ".gamma"] = 1.0 + torch.randn(*shape, generator=g) * gamma_std

    def resblock(p, i, o):
        gamma(p + "norm1", i, False); conv3(p + "conv1", o, i, (3, 3, 3))
        gamma(p + "norm2", o, False); conv3(p + "conv2", o, o, (3, 3, 3))
        if i != o:
            conv3(p + "conv_shortcut", o, i, (1, 1, 1))

    z = cfg.z_dim
    conv3("post_quant_conv", z, z, (1, 1, 1))
    d0, blocks = cfg.decoder_dims()
    conv3("decoder.conv_in", d0, z, (3, 3, 3))
    resblock("decoder.mid_block.resnets.0.", d0, d0)
    gamma("decoder.mid_block.attentions.0.norm", d0, True)
    conv2("decoder.mid_block.attentions.0.to_qkv", 3 * d0, d0, 1)
    conv2("decoder.mid_block.attentions.0.proj", d0, d0, 1)
    resblock("decoder.mid_block.resnets.1.", d0, d0)
    last = d0
    for bi, (i, o, mode) in enumerate(blocks):
        cur = i
        for j in range(cfg.num_res_blocks + 1):
            resblock(f"decoder.up_blocks.{bi}.resnets.{j}.", cur, o)
            cur = o
        if mode is not None:
            conv2(f"decoder.up_blocks.{bi}.upsamplers.0.resample.1", o // 2, o, 3)
            if mode == "upsample3d":
                conv3(f"decoder.up_blocks.{bi}.upsamplers.0.time_conv", 2 * o, o, (3, 1, 1))
        last = o
    gamma("decoder.norm_out", last, False)
    conv3("decoder.conv_out", 3, last, (3, 3, 3))
    if with_encoder:
        # QwenImageEncoder3d: down_blocks is ONE flat ModuleList (residual blocks and resamples interleaved)
        stages = encoder_dims(cfg)
        conv3("encoder.conv_in", stages[0][0], 3, (3, 3, 3))
        k = 0
        for (i, o, mode) in stages:
            cur = i
            for _ in range(cfg.num_res_blocks):
                resblock(f"encoder.down_blocks.{k}.", cur, o)
                cur = o
                k += 1
            if mode is not None:
                conv2(f"encoder.down_blocks.{k}.resample.1", o, o, 3)
                if mode == "downsample3d":
                    conv3(f"encoder.down_blocks.{k}.time_conv", o, o, (3, 1, 1))
                k += 1
        top = stages[-1][1]
        resblock("encoder.mid_block.resnets.0.", top, top)
        gamma("encoder.mid_block.attentions.0.norm", top, True)
        conv2("encoder.mid_block.attentions.0.to_qkv", 3 * top, top, 1)
        conv2("encoder.mid_block.attentions.0.proj", top, top, 1)
        resblock("encoder.mid_block.resnets.1.", top, top)
        gamma("encoder.norm_out", top, False)
        conv3("encoder.conv_out", 2 * z, top, (3, 3, 3))
        conv3("quant_conv", 2 * z, 2 * z, (1, 1, 1))
    return sd


# ------------------------------------------------------------------ building blocks
def causal_conv3d(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """QwenImageCausalConv3d with padding = (k-1)/2 per axis: spatial padding symmetric, time padding
    2*p in front only (the frame never sees the future)."""
    kt, kh, kw = w.shape[2:]
    pt, ph, pw = (kt - 1) // 2, (kh - 1) // 2, (kw - 1) // 2
    x = F.pad(x, (pw, pw, ph, ph, 2 * pt, 0))
    return F.conv3d(x, w, b)


def rms_norm_channels(x: torch.Tensor, gamma: torch.Tensor) -> torch.Tensor:
    """QwenImageRMS_norm(channel_first=True): L2-normalise over channels, times sqrt(C), times gamma."""
    return F.normalize(x, dim=1) * (x.shape[1] ** 0.5) * gamma


def residual_block(sd, p: str, x: torch.Tensor) -> torch.Tensor:
    h = causal_conv3d(x, sd[p + "conv_shortcut.weight"], sd[p + "conv_shortcut.bias"]) \
        if (p + "conv_shortcut.weight") in sd else x
    x = F.silu(rms_norm_channels(x, sd[p + "norm1.gamma"]))
    x = causal_conv3d(x, sd[p + "conv1.weight"], sd[p + "conv1.bias"])
    x = F.silu(rms_norm_channels(x, sd[p + "norm2.gamma"]))
    x = causal_conv3d(x, sd[p + "conv2.weight"], sd[p + "conv2.bias"])
    return x + h


def attention_block(sd, p: str, x: torch.Tensor) -> torch.Tensor:
    """Single-head self-attention over the H*W positions of every frame."""
    identity = x
    b, c, t, h, w = x.shape
    x = x.permute(0, 2, 1, 3, 4).reshape(b * t, c, h, w)
    x = rms_norm_channels(x, sd[p + "norm.gamma"])
    qkv = F.conv2d(x, sd[p + "to_qkv.weight"], sd[p + "to_qkv.bias"])
    qkv = qkv.reshape(b * t, 1, c * 3, -1).permute(0, 1, 3, 2).contiguous()
    q, k, v = qkv.chunk(3, dim=-1)
    x = F.scaled_dot_product_attention(q, k, v)
    x = x.squeeze(1).permute(0, 2, 1).reshape(b * t, c, h, w)
    x = F.conv2d(x, sd[p + "proj.weight"], sd[p + "proj.bias"])
    x = x.view(b, t, c, h, w).permute(0, 2, 1, 3, 4)
    return x + identity


def resample_up(sd, p: str, x: torch.Tensor) -> torch.Tensor:
    """QwenImageResample 'upsample2d'/'upsample3d' on the FIRST chunk of a clip: the temporal conv is
    skipped (upstream marks the cache slot 'Rep' and only upsamples spatially), so both modes are
    nearest-exact 2x in fp32 followed by a 3x3 conv that halves the channels."""
    b, c, t, h, w = x.shape
    x = x.permute(0, 2, 1, 3, 4).reshape(b * t, c, h, w)
    x = F.interpolate(x.float(), scale_factor=(2.0, 2.0), mode="nearest-exact").type_as(x)
    x = F.conv2d(x, sd[p + "resample.1.weight"], sd[p + "resample.1.bias"], padding=1)
    return x.view(b, t, x.shape[1], x.shape[2], x.shape[3]).permute(0, 2, 1, 3, 4)


def decode(sd, cfg: VAEOracleConfig, z: torch.Tensor) -> torch.Tensor:
    """AutoencoderKLQwenImage.decode on a one-frame latent [B, z, 1, H, W] -> [B, 3, 1, 8H, 8W] in
    [-1, 1].  (Clips with T > 1 need the feature cache of upstream's chunked decode; the image pipeline
    never produces them.)"""
    if z.dim() == 4:
        z = z.unsqueeze(2)
    if z.shape[2] != 1:
        raise ValueError("the image pipeline decodes exactly one latent frame")
    x = causal_conv3d(z, sd["post_quant_conv.weight"], sd["post_quant_conv.bias"])
    x = causal_conv3d(x, sd["decoder.conv_in.weight"], sd["decoder.conv_in.bias"])
    x = residual_block(sd, "decoder.mid_block.resnets.0.", x)
    x = attention_block(sd, "decoder.mid_block.attentions.0.", x)
    x = residual_block(sd, "decoder.mid_block.resnets.1.", x)
    _, blocks = cfg.decoder_dims()
    for bi, (_, _, mode) in enumerate(blocks):
        for j in range(cfg.num_res_blocks + 1):
            x = residual_block(sd, f"decoder.up_blocks.{bi}.resnets.{j}.", x)
        if mode is not None:
            x = resample_up(sd, f"decoder.up_blocks.{bi}.upsamplers.0.", x)
    x = F.silu(rms_norm_channels(x, sd["decoder.norm_out.gamma"]))
    x = causal_conv3d(x, sd["decoder.conv_out.weight"], sd["decoder.conv_out.bias"])
    return torch.clamp(x, min=-1.0, max=1.0)


def resample_down(sd, p: str, x: torch.Tensor) -> torch.Tensor:
    """QwenImageResample 'downsample2d'/'downsample3d' on the FIRST chunk of a clip: per frame ZeroPad2d((0,1,0,1))
    then a 3x3 stride-2 conv; the temporal stride-2 conv of 'downsample3d' only runs from the second chunk on
    (upstream stores the first chunk in its feature cache and returns), so an image never meets it."""
    b, c, t, h, w = x.shape
    x = x.permute(0, 2, 1, 3, 4).reshape(b * t, c, h, w)
    x = F.pad(x, (0, 1, 0, 1))
    x = F.conv2d(x, sd[p + "resample.1.weight"], sd[p + "resample.1.bias"], stride=2)
    return x.view(b, t, x.shape[1], x.shape[2], x.shape[3]).permute(0, 2, 1, 3, 4)


def encode(sd, cfg: VAEOracleConfig, x: torch.Tensor) -> torch.Tensor:
    """AutoencoderKLQwenImage._encode on ONE frame [B, 3, 1, H, W] in [-1, 1] -> posterior moments
    [B, 2*z, 1, H/8, W/8] (mean | logvar): encoder (conv_in, flat down_blocks, mid block, norm/SiLU/conv_out) then
    quant_conv.  H and W must be multiples of 8."""
    if x.dim() == 4:
        x = x.unsqueeze(2)
    if x.shape[2] != 1:
        raise ValueError("the image pipeline encodes exactly one frame")
    x = causal_conv3d(x, sd["encoder.conv_in.weight"], sd["encoder.conv_in.bias"])
    k = 0
    for (_, _, mode) in encoder_dims(cfg):
        for _ in range(cfg.num_res_blocks):
            x = residual_block(sd, f"encoder.down_blocks.{k}.", x)
            k += 1
        if mode is not None:
            x = resample_down(sd, f"encoder.down_blocks.{k}.", x)
            k += 1
    x = residual_block(sd, "encoder.mid_block.resnets.0.", x)
    x = attention_block(sd, "encoder.mid_block.attentions.0.", x)
    x = residual_block(sd, "encoder.mid_block.resnets.1.", x)
    x = F.silu(rms_norm_channels(x, sd["encoder.norm_out.gamma"]))
    x = causal_conv3d(x, sd["encoder.conv_out.weight"], sd["encoder.conv_out.bias"])
    return causal_conv3d(x, sd["quant_conv.weight"], sd["quant_conv.bias"])


def encode_mode(sd, cfg: VAEOracleConfig, x: torch.Tensor) -> torch.Tensor:
    """``retrieve_latents(vae.encode(image), sample_mode="argmax")``: the posterior's mode = its mean."""
    return encode(sd, cfg, x)[:, :cfg.z_dim]


def normalize_latents(latents: torch.Tensor, cfg: VAEOracleConfig) -> torch.Tensor:
    """QwenImageEditPipeline._encode_vae_image after the encode: ``(latents - mean) / std`` per latent channel."""
    shape = [1, cfg.z_dim] + [1] * (latents.dim() - 2)
    mean = torch.tensor(cfg.latents_mean).view(shape).to(latents.device, latents.dtype)
    std = torch.tensor(cfg.latents_std).view(shape).to(latents.device, latents.dtype)
    return (latents - mean) / std


def calculate_dimensions(target_area: int, ratio: float):
    """QwenImageEditPipeline's output size for an input image of aspect ratio width/height: the area is
    target_area, both sides rounded to multiples of 32.  Returns (width, height)."""
    import math
    width = math.sqrt(target_area * ratio)
    height = width / ratio
    return int(round(width / 32) * 32), int(round(height / 32) * 32)


def denormalize_latents(latents: torch.Tensor, cfg: VAEOracleConfig) -> torch.Tensor:
    """QwenImagePipeline.__call__ before vae.decode: ``latents / (1/std) + mean`` per latent channel."""
    mean = torch.tensor(cfg.latents_mean).view(1, cfg.z_dim, 1, 1, 1).to(latents.device, latents.dtype)
    inv_std = 1.0 / torch.tensor(cfg.latents_std).view(1, cfg.z_dim, 1, 1, 1).to(latents.device, latents.dtype)
    return latents / inv_std + mean


def decode_packed(sd, cfg: VAEOracleConfig, tokens: torch.Tensor, Hl: int, Wl: int) -> torch.Tensor:
    """packed latents [B, S, 64] -> image [B, 3, 8*Hl, 8*Wl] (unpack, de-normalise, decode, frame 0)."""
    B, _, ch = tokens.shape
    x = tokens.view(B, Hl // 2, Wl // 2, ch // 4, 2, 2).permute(0, 3, 1, 4, 2, 5).reshape(B, ch // 4, 1, Hl, Wl)
    return decode(sd, cfg, denormalize_latents(x, cfg))[:, :, 0]


def psnr(a: torch.Tensor, b: torch.Tensor, data_range: float = 2.0) -> float:
    """Peak signal-to-noise ratio in dB for images in [-1, 1]."""
    mse = float((a.double() - b.double()).pow(2).mean())
    if mse == 0:
        return float("inf")
    import math
    return 10.0 * math.log10(data_range ** 2 / mse)
