"""Drop-in for the reference's vLLM-facing model wrapper.

Mirrors ``QwenImageGenModel`` (reference src/qwen_image_gen/model.py:17-165) and ``QwenImageOutputs``
(src/qwen_image_gen/types.py:35-39): same names, argument meaning and error behaviour, with the B200
transformer behind them.  The reference's ``QwenImageConfig`` / ``create_qwen_image_config`` (config.py) are
out of scope (the model reads ``model_id`` / subfolders off ``vllm_config.model_config.hf_config``
with the same defaults, and this package's own shapes live in ``config.py``).  Host layer only — the loop body is the reference's own three tensor ops
(``flatten/transpose``, ``transpose/unflatten``, ``latents - residual * 0.1``, model.py:117-135); every
FLOP of ``self.transformer(...)`` runs in libqi_b200.so.

Differences, all host-side and stated here:
* the components are injected (``transformer=``, ``vae=``) or loaded from a *local* directory — there is
  no hub access; a load failure raises the reference's ``RuntimeError("Failed to load QwenImage
  components: ...")`` (model.py:72-75);
* the modulation of every step is hoisted before the loop (``prepare_schedule``): the timesteps are
  known up front, so the 2L+1 modulation Linears run as one tcgen05 GEMM instead of once per step;
* ``timesteps`` must be 1-D ``[steps]``.  The reference iterates ``enumerate(timesteps)`` and calls
  ``t.expand(batch_size)`` (model.py:117-119), which only works for 1-D input; a 2-D tensor raises
  ``ValueError`` here instead of the reference's shape error from ``expand``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Optional

import torch
import torch.nn as nn

from .adapter import QwenImageInputs, QwenImageOutputMode, QwenImageTask  # noqa: F401  (re-exported)
from .transformer import B200QwenImageTransformer2DModel


@dataclass
class QwenImageOutputs:
    """src/qwen_image_gen/types.py:35-39."""
    latents: torch.Tensor
    pixels: Optional[torch.Tensor] = None


class QwenImageGenModel(nn.Module):
    """vLLM-facing wrapper: ``forward(**kwargs)`` pulls ``kwargs["qwen_inputs"]`` and runs the
    reference's denoising loop on the B200 transformer (model.py:77-143)."""

    def __init__(self, vllm_config: Any = None, prefix: str = "", *, transformer=None, vae=None,
                 device=None, dtype: Optional[torch.dtype] = None):
        super().__init__()
        self.vllm_config = vllm_config
        model_config = getattr(vllm_config, "model_config", None)
        self.config = getattr(model_config, "hf_config", None)
        self.dtype = dtype if dtype is not None else getattr(model_config, "dtype", torch.bfloat16)
        self.device = torch.device(device) if device is not None else torch.device(
            "cuda" if torch.cuda.is_available() else "cpu")
        if transformer is not None:
            self.transformer, self.vae = transformer, vae
            self.device = transformer.device
        else:
            self._init_qwen_image_components()

    def _init_qwen_image_components(self) -> None:
        model_id = getattr(self.config, "model_id", "Qwen/Qwen-Image")
        transformer_subfolder = getattr(self.config, "transformer_subfolder", "transformer")
        vae_subfolder = getattr(self.config, "vae_subfolder", "vae")
        try:
            self.transformer = B200QwenImageTransformer2DModel.from_pretrained(
                model_id, subfolder=transformer_subfolder, torch_dtype=torch.bfloat16, device=self.device)
            from .vae import B200QwenImageVAE
            self.vae = B200QwenImageVAE.from_pretrained(model_id, subfolder=vae_subfolder, device=self.device)
        except ImportError:
            raise
        except Exception as e:
            raise RuntimeError(f"Failed to load QwenImage components: {e}") from e

    def forward(self, input_ids: Optional[torch.Tensor] = None, positions: Optional[torch.Tensor] = None,
                inputs_embeds: Optional[torch.Tensor] = None, **kwargs: Any) -> torch.Tensor:
        qwen_inputs = kwargs.get("qwen_inputs")
        if qwen_inputs is None:
            # dummy output for vLLM's profile / warm-up runs (model.py:88-90)
            return torch.empty(0, device=self.device, dtype=self.dtype)
        if not isinstance(qwen_inputs, QwenImageInputs):
            raise ValueError("qwen_inputs must be of type QwenImageInputs")
        return self._generate_image(qwen_inputs)

    @torch.no_grad()
    def _generate_image(self, inputs: QwenImageInputs) -> torch.Tensor:
        tr = self.transformer
        dev, dt = tr.device, tr.dtype
        prompt_embeds = inputs.prompt_embeds.to(device=dev, dtype=dt)
        prompt_mask = inputs.prompt_embeds_mask.to(device=dev)
        latents = inputs.image_latents.to(device=dev, dtype=dt)
        timesteps = inputs.timesteps.to(device=dev)
        if timesteps.dim() != 1:
            raise ValueError("timesteps must be 1-D [steps] (each entry is expanded to the batch)")
        if latents.dim() not in (4, 5):
            raise ValueError("image_latents must be 4D or 5D")
        batch_size = latents.shape[0]
        guidance = None
        if inputs.guidance_scale > 1.0:
            guidance = torch.tensor([inputs.guidance_scale] * batch_size, device=dev, dtype=dt)
        # temb depends only on t: all modulation vectors of the run in one GEMM before the loop
        tr.prepare_schedule(timesteps.to(torch.float32))
        try:
            for i, t in enumerate(timesteps):
                out = tr(hidden_states=latents.flatten(2).transpose(1, 2), encoder_hidden_states=prompt_embeds,
                         encoder_hidden_states_mask=prompt_mask, timestep=t.expand(batch_size), guidance=guidance,
                         return_dict=True, _step_index=i)
                residual = out.sample.transpose(1, 2).unflatten(2, latents.shape[2:])
                latents = latents - residual * 0.1
        finally:
            tr.clear_schedule()
        if inputs.output_mode.value == "pixels":
            if self.vae is None:
                raise RuntimeError("output_mode requests pixels but no VAE was provided")
            pixels = self.vae.decode(latents)
            return getattr(pixels, "sample", pixels)
        return latents

    # vLLM model-interface placeholders (model.py:145-165)
    def get_multimodal_embeddings(self, **kwargs: Any) -> torch.Tensor:
        return torch.empty(0, device=self.device, dtype=self.dtype)

    def compute_logits(self, hidden_states: torch.Tensor, sampling_metadata: Any = None) -> Optional[torch.Tensor]:
        return None

    def load_weights(self, weights: Any) -> set:
        return set()

    def get_mm_mapping(self) -> dict:
        return {}
