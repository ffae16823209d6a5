"""Prompt-embedding producer — host layer.

The denoising path consumes ``prompt_embeds [B, T, 3584]`` + mask.  The reference gets them from
Qwen2.5-VL hidden states through diffusers' ``get_qwen_prompt_embeds`` (call site
src/qwen_image/processor.py:52-60: chat template, then the first ``start_idx = 34`` template tokens are
dropped) or, inside the runner adapter, from a bare tokenizer + encoder pass without template
(src/qwen_image/runner_adapter.py:70-108).  Both are restated here over *injected* ``text_encoder`` /
``tokenizer`` objects (any HF-style callables: the encoder itself is a library model and out of scope).
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple, Union

import torch

PROMPT_TEMPLATE_ENCODE = (
    "<|im_start|>system\nDescribe the image by detailing the color, "
    "shape, size, texture, quantity, text, spatial relationships of the "
    "objects and background:<|im_end|>\n<|im_start|>user\n{}<|im_end|>\n"
    "<|im_start|>assistant\n"
)
PROMPT_TEMPLATE_ENCODE_START_IDX = 34


def _valid_rows(hidden_states: torch.Tensor, mask: torch.Tensor) -> List[torch.Tensor]:
    """Per sample, the hidden states at the positions the attention mask marks valid (any padding side)."""
    keep = mask.bool()
    lengths = keep.sum(dim=1).tolist()
    return list(torch.split(hidden_states[keep], lengths, dim=0))


def get_qwen_prompt_embeds(text_encoder, tokenizer, prompt: Union[str, Sequence[str]],
                           prompt_template_encode: str = PROMPT_TEMPLATE_ENCODE,
                           prompt_template_encode_start_idx: int = PROMPT_TEMPLATE_ENCODE_START_IDX,
                           tokenizer_max_length: int = 1024, device=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """-> (prompt_embeds [B, T, D], mask [B, T] int64): wrap every prompt in the chat template, encode,
    take the last hidden layer, strip padding, drop the template prefix, right-pad to the longest."""
    prompts = [prompt] if isinstance(prompt, str) else list(prompt)
    drop = int(prompt_template_encode_start_idx)
    text = [prompt_template_encode.format(p) for p in prompts]
    tokens = tokenizer(text, max_length=tokenizer_max_length + drop, padding=True, truncation=True,
                       return_tensors="pt")
    if device is not None:
        tokens = tokens.to(device)
    out = text_encoder(input_ids=tokens.input_ids, attention_mask=tokens.attention_mask,
                       output_hidden_states=True)
    hidden = out.hidden_states[-1]
    rows = [r[drop:] for r in _valid_rows(hidden, tokens.attention_mask)]
    longest = max(r.size(0) for r in rows)
    embeds = torch.stack([torch.cat([r, r.new_zeros(longest - r.size(0), r.size(1))]) for r in rows])
    mask = torch.stack([torch.cat([torch.ones(r.size(0), dtype=torch.long, device=r.device),
                                   torch.zeros(longest - r.size(0), dtype=torch.long, device=r.device)])
                        for r in rows])
    return embeds, mask


def encode_prompt_inline(worker, raw_prompt: Optional[str], device, raw_image=None, edit: bool = False):
    """The runner adapter's on-the-fly path (runner_adapter.py:70-108), literally: no template, no prefix
    drop; last hidden layer and the tokenizer's attention mask.  Raises the reference's ValueErrors."""
    if worker is None or not hasattr(worker, "qwen_text_encoder") or worker.qwen_text_encoder is None:
        raise ValueError("prompt_embeds missing and Qwen2.5-VL not available on worker")
    if raw_prompt is None:
        raise ValueError("raw_prompt required to compute embeddings on the fly")
    if edit:
        if getattr(worker, "qwen_processor", None) is None or raw_image is None:
            raise ValueError("processor and raw_image required for edit embeddings")
        proc = worker.qwen_processor(text=[raw_prompt], images=raw_image, padding=True, return_tensors="pt").to(device)
        out = worker.qwen_text_encoder(input_ids=proc.input_ids, attention_mask=proc.attention_mask,
                                       pixel_values=proc.pixel_values, image_grid_thw=proc.image_grid_thw,
                                       output_hidden_states=True)
        return out.hidden_states[-1], proc.attention_mask
    tok = worker.qwen_tokenizer([raw_prompt], padding=True, truncation=True, return_tensors="pt").to(device)
    out = worker.qwen_text_encoder(input_ids=tok.input_ids, attention_mask=tok.attention_mask,
                                   output_hidden_states=True)
    return out.hidden_states[-1], tok.attention_mask


def build_custom_inputs_text_only(*, text_encoder, tokenizer, image_latents: torch.Tensor, device, dtype,
                                  prompt: str, tokenizer_max_length: int = 1024,
                                  prompt_template_encode: str = PROMPT_TEMPLATE_ENCODE,
                                  prompt_template_encode_start_idx: int = PROMPT_TEMPLATE_ENCODE_START_IDX,
                                  num_inference_steps: int = 30, guidance_scale: float = 4.0,
                                  seed: Optional[int] = None, img_shapes=None, txt_seq_lens=None):
    """src/qwen_image/processor.py:20-106 with one stated difference: the reference VAE-*encodes* a pixel
    ``image`` into the initial latents with diffusers' modular-pipeline steps; the VAE encoder is outside
    this path, so the caller passes ``image_latents`` directly."""
    from .adapter import QwenImageCustomInputs, validate_custom_inputs
    embeds, mask = get_qwen_prompt_embeds(text_encoder, tokenizer, prompt=prompt,
                                          prompt_template_encode=prompt_template_encode,
                                          prompt_template_encode_start_idx=prompt_template_encode_start_idx,
                                          tokenizer_max_length=tokenizer_max_length, device=device)
    custom = QwenImageCustomInputs(prompt_embeds=embeds, prompt_embeds_mask=mask,
                                   image_latents=image_latents.to(device=device, dtype=dtype),
                                   img_shapes=img_shapes, txt_seq_lens=txt_seq_lens,
                                   num_inference_steps=num_inference_steps, guidance_scale=guidance_scale, seed=seed)
    validate_custom_inputs(custom)
    return custom
