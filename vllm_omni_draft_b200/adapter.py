"""Boundary #3: the thin Python façade the reference's vLLM worker calls (host layer).

Mirrors ``QwenImageRunnerAdapter`` / ``QwenImageCustomInputs`` / ``validate_custom_inputs``
(reference src/qwen_image/runner_adapter.py:24-160, src/qwen_image/types.py:29-106) and the
simplified ``QwenImageInputs`` bundle (src/qwen_image_gen/types.py:22-32): same names, argument
meaning and error behaviour, with the B200 transformer behind it.  The reference's own loop is a
placeholder ("simple linear schedule placeholder", runner_adapter.py:62-68; ``latents - residual``,
:141); ``schedule="reference_placeholder"`` reproduces it literally, the default runs the real
FlowMatch-Euler / true-CFG loop of the pipeline the reference benchmarks."""
from __future__ import annotations

from dataclasses import dataclass, field
from enum import Enum
from typing import List, Optional, Sequence, Tuple

import torch

from .pipeline import B200QwenImagePipeline, pack_latents, unpack_latents

TASK_IMAGE_GENERATION: str = "IMAGE_GENERATION"
CUSTOM_INPUTS_KEY: str = "qwen_image_custom_inputs"


@dataclass
class QwenImageCustomInputs:
    """Request-side tensor bundle (field-for-field the reference's schema, types.py:51-78)."""
    prompt_embeds: torch.Tensor
    prompt_embeds_mask: torch.Tensor
    image_latents: torch.Tensor
    control_image_latents: Optional[List[torch.Tensor]] = None
    img_shapes: Optional[Sequence[Tuple[int, int, int]]] = None
    txt_seq_lens: Optional[Sequence[int]] = None

    class QwenImageTask(str, Enum):
        T2I = "T2I"
        I2I = "I2I"
        TI2I = "TI2I"

    class QwenImageOutputMode(str, Enum):
        PIXELS = "PIXELS"
        LATENTS = "LATENTS"
        PIXELS_AND_LATENTS = "PIXELS_AND_LATENTS"
        PIXELS_AND_MASK = "PIXELS_AND_MASK"

    task: QwenImageTask = QwenImageTask.T2I
    output_mode: QwenImageOutputMode = QwenImageOutputMode.PIXELS
    num_inference_steps: int = 30
    guidance_scale: float = 4.0
    seed: Optional[int] = None
    height: Optional[int] = None
    width: Optional[int] = None
    extras: dict = field(default_factory=dict)


def validate_custom_inputs(inputs: QwenImageCustomInputs) -> None:
    """Same checks and messages as the reference (types.py:84-106); ValueError on any violation."""
    if inputs.prompt_embeds.dim() != 3:
        raise ValueError("prompt_embeds must be 3D [B, T_txt, D_txt]")
    if inputs.prompt_embeds_mask.dim() != 2:
        raise ValueError("prompt_embeds_mask must be 2D [B, T_txt]")
    if inputs.prompt_embeds.shape[0] != inputs.prompt_embeds_mask.shape[0]:
        raise ValueError("Batch size mismatch between embeds and mask")
    if inputs.image_latents.dim() not in (4, 5):
        raise ValueError("image_latents must be 4D or 5D (variant dependent)")
    if inputs.txt_seq_lens is not None and inputs.img_shapes is not None:
        if len(inputs.txt_seq_lens) != len(inputs.img_shapes):
            raise ValueError("txt_seq_lens and img_shapes length mismatch")


class QwenImageTask(str, Enum):
    TEXT_TO_IMAGE = "text_to_image"
    IMAGE_TO_IMAGE = "image_to_image"


class QwenImageOutputMode(str, Enum):
    LATENTS = "latents"
    PIXELS = "pixels"


@dataclass
class QwenImageInputs:
    """src/qwen_image_gen/types.py:22-32."""
    prompt_embeds: torch.Tensor
    prompt_embeds_mask: torch.Tensor
    image_latents: torch.Tensor
    timesteps: torch.Tensor
    guidance_scale: float = 4.0
    num_inference_steps: int = 30
    task: QwenImageTask = QwenImageTask.TEXT_TO_IMAGE
    output_mode: QwenImageOutputMode = QwenImageOutputMode.PIXELS


class QwenImageRunnerAdapter:
    def __init__(self, *, transformer, vae=None, device=None, dtype=torch.bfloat16, worker: Optional[object] = None):
        self.transformer = transformer
        self.vae = vae
        self.device = device if device is not None else transformer.device
        self.dtype = dtype
        self.worker = worker
        self._pipe = B200QwenImagePipeline(transformer, vae=vae)

    @torch.inference_mode()
    def generate(self, inputs: QwenImageCustomInputs, *, decode_pixels: bool = True,
                 schedule: str = "flowmatch", negative_prompt_embeds: Optional[torch.Tensor] = None) -> torch.Tensor:
        if inputs.prompt_embeds is None:
            # embeddings on the fly when the worker holds a Qwen2.5-VL encoder (runner_adapter.py:70-108);
            # same ValueErrors as the reference when it does not
            from .prompt import encode_prompt_inline
            hidden, mask = encode_prompt_inline(
                self.worker, getattr(inputs, "raw_prompt", None), self.device,
                raw_image=getattr(inputs, "raw_image", None),
                edit=inputs.task == QwenImageCustomInputs.QwenImageTask.TI2I)
            inputs.prompt_embeds, inputs.prompt_embeds_mask = hidden, mask
        validate_custom_inputs(inputs)
        latents = inputs.image_latents.to(device=self.device, dtype=self.dtype)
        bsz = latents.shape[0]
        num_steps = max(int(inputs.num_inference_steps), 1)
        prompt_embeds = inputs.prompt_embeds.to(device=self.device, dtype=self.dtype)
        prompt_mask = inputs.prompt_embeds_mask.to(device=self.device)
        Mode = QwenImageCustomInputs.QwenImageOutputMode

        if schedule == "reference_placeholder":
            # literal restatement of runner_adapter.py:64-68, 118-141 (needs 64-channel token latents)
            timesteps = (torch.linspace(1.0, 0.0, steps=num_steps, device=self.device) * 1000).to(torch.long)
            guidance = None
            if inputs.guidance_scale is not None and inputs.guidance_scale > 1.0:
                guidance = torch.tensor([inputs.guidance_scale] * bsz, device=self.device, dtype=self.dtype)
            for t in timesteps:
                out = self.transformer(hidden_states=latents.flatten(2).transpose(1, 2),
                                       encoder_hidden_states=prompt_embeds, encoder_hidden_states_mask=prompt_mask,
                                       timestep=t.expand(bsz), img_shapes=inputs.img_shapes,
                                       txt_seq_lens=inputs.txt_seq_lens, guidance=guidance, return_dict=True)
                residual = out.sample.transpose(1, 2).unflatten(2, latents.shape[2:])
                latents = latents - residual
        elif schedule == "flowmatch":
            lat4 = latents[:, :, 0] if latents.dim() == 5 else latents
            Hl, Wl = lat4.shape[-2:]
            scale = float(inputs.guidance_scale or 1.0)
            packed = self._pipe(prompt_embeds=prompt_embeds, prompt_embeds_mask=prompt_mask,
                                negative_prompt_embeds=negative_prompt_embeds, true_cfg_scale=scale,
                                num_inference_steps=num_steps, height=Hl * 8, width=Wl * 8,
                                latents=pack_latents(lat4), output_type="latent").images
            latents = unpack_latents(packed, Hl, Wl)
            if inputs.image_latents.dim() == 4:
                latents = latents[:, :, 0]
        else:
            raise ValueError(f"unknown schedule {schedule!r}")

        if inputs.output_mode == Mode.LATENTS or not decode_pixels:
            return latents
        if self.vae is None:
            raise RuntimeError("output_mode requests pixels but no VAE was provided")
        if schedule == "flowmatch":
            # the real pipeline de-normalises with the VAE's channel statistics before decoding
            lat5 = latents if latents.dim() == 5 else latents.unsqueeze(2)
            return self._pipe.decode_latents(pack_latents(lat5[:, :, 0]), lat5.shape[-2], lat5.shape[-1])
        pixels = self.vae.decode(latents)        # literal: runner_adapter.py:149-160
        return getattr(pixels, "sample", pixels)
