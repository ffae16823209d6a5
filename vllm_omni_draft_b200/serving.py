"""vLLM V1 wiring for the B200 path — host layer, pure Python.

Mirrors the reference's two generations of glue:

* ``qwen_image/v1``: ``QwenImageWorker`` (src/qwen_image/v1/worker.py:18-82 — builds a
  ``QwenImageRunnerAdapter`` from ``model_runner.qwen_transformer`` / ``qwen_vae``, optionally loads a
  Qwen2.5-VL encoder when ``QWEN_VL_ENABLE=1``) and ``QwenImageUniProcExecutor``
  (src/qwen_image/v1/executor.py:16-31 — injects the worker class unless the user set one);
* ``qwen_image_gen``: ``QwenImageGenWorker`` (src/qwen_image_gen/worker.py:16-110 — detects requests whose
  multimodal inputs carry ``processing_info.extras["qwen_inputs"]`` and runs ``model._generate_image``) and
  ``QwenImageGenExecutor`` (src/qwen_image_gen/executor.py:11-31).

The reference's modules subclass vLLM classes at import time and are broken against the installed vLLM
(``vllm.vllm.v1...`` import path, a ``scheduled_multimodal_inputs`` field SchedulerOutput does not have).  Here
the request-handling logic lives in mixins that need nothing from vLLM (so it is testable and reusable), and
``make_*`` factories bind them onto vLLM's ``Worker`` / ``UniProcExecutor`` lazily when vLLM is importable.
"""
from __future__ import annotations

import os
from typing import Any, Optional

import torch

from .adapter import QwenImageInputs, QwenImageRunnerAdapter

QWEN_IMAGE_WORKER_CLS = "vllm_omni_draft_b200.serving.QwenImageWorker"
QWEN_IMAGE_GEN_WORKER_CLS = "vllm_omni_draft_b200.serving.QwenImageGenWorker"


# ------------------------------------------------------------------ executor side
def inject_worker_cls(vllm_config: Any, worker_cls: Any) -> Any:
    """``if not parallel_config.worker_cls: parallel_config.worker_cls = ours`` — an explicit user choice is
    respected (qwen_image/v1/executor.py:26-31, qwen_image_gen/executor.py:14-27)."""
    pc = vllm_config.parallel_config
    if not getattr(pc, "worker_cls", None) or pc.worker_cls == "auto":
        pc.worker_cls = worker_cls
    return vllm_config


# ------------------------------------------------------------------ worker side: qwen_image/v1
class QwenImageWorkerMixin:
    """Adapter plumbing of ``QwenImageWorker``; expects ``self.model_runner`` and ``self.device``."""

    _qwen_image_adapter: Optional[QwenImageRunnerAdapter] = None
    qwen_text_encoder = qwen_tokenizer = qwen_processor = None

    def _load_qwen_vl(self) -> None:
        """Optional Qwen2.5-VL for on-the-fly prompt embeddings; fails soft like the reference (worker.py:27-55)."""
        if os.getenv("QWEN_VL_ENABLE", "0") != "1":
            return
        model_id = os.getenv("QWEN_VL_MODEL_ID", "Qwen/Qwen2.5-VL-7B-Instruct")
        dtype_str = os.getenv("QWEN_VL_DTYPE", "auto")
        model = getattr(getattr(self, "model_runner", None), "model", None)
        q_dtype = getattr(model, "dtype", torch.float16) if dtype_str == "auto" else getattr(torch, dtype_str)
        try:
            from transformers import AutoProcessor, AutoTokenizer, Qwen2_5_VLForConditionalGeneration
            index = getattr(self.device, "index", None) or 0
            self.qwen_text_encoder = Qwen2_5_VLForConditionalGeneration.from_pretrained(
                model_id, torch_dtype=q_dtype, device_map={"": index}).eval()
            self.qwen_tokenizer = AutoTokenizer.from_pretrained(model_id)
            try:
                self.qwen_processor = AutoProcessor.from_pretrained(model_id)
            except Exception:
                self.qwen_processor = None
        except Exception:
            self.qwen_text_encoder = self.qwen_tokenizer = self.qwen_processor = None

    def get_qwen_image_adapter(self) -> Optional[QwenImageRunnerAdapter]:
        if self._qwen_image_adapter is not None:
            return self._qwen_image_adapter
        runner = getattr(self, "model_runner", None)
        transformer = getattr(runner, "qwen_transformer", None)
        vae = getattr(runner, "qwen_vae", None)
        if transformer is None or vae is None:
            return None
        dtype = getattr(getattr(runner, "model", None), "dtype", torch.bfloat16)
        self._qwen_image_adapter = QwenImageRunnerAdapter(transformer=transformer, vae=vae, device=self.device,
                                                          dtype=dtype, worker=self)
        return self._qwen_image_adapter


# ------------------------------------------------------------------ worker side: qwen_image_gen
def _qwen_inputs_of(mm_input: Any) -> Optional[QwenImageInputs]:
    info = getattr(mm_input, "processing_info", None)
    extras = getattr(info, "extras", None) if info else None
    if extras and "qwen_inputs" in extras:
        return extras["qwen_inputs"]
    return None


class QwenImageGenWorkerMixin:
    """Request detection + dispatch of ``QwenImageGenWorker`` (worker.py:47-102)."""

    def verify_qwen_components(self) -> None:
        model = getattr(getattr(self, "model_runner", None), "model", None)
        if model is not None and not (hasattr(model, "transformer") and hasattr(model, "vae")):
            raise RuntimeError("QwenImage components not loaded properly. transformer and vae are required.")

    def _is_image_generation_request(self, scheduler_output: Any) -> bool:
        inputs = getattr(scheduler_output, "scheduled_multimodal_inputs", None)
        if not inputs:
            return False
        return any(_qwen_inputs_of(m) is not None for m in inputs)

    def _execute_image_generation(self, scheduler_output: Any) -> Optional[Any]:
        model = self.model_runner.model
        for mm_input in scheduler_output.scheduled_multimodal_inputs:
            qwen_inputs = _qwen_inputs_of(mm_input)
            if qwen_inputs is None:
                continue
            with torch.no_grad():
                generated = model._generate_image(qwen_inputs)
            mm_input.generated_output = generated      # one request per execute_model, like the reference
            return generated
        return None

    def execute_model(self, scheduler_output: Any, *args: Any, **kwargs: Any) -> Optional[Any]:
        if self._is_image_generation_request(scheduler_output):
            return self._execute_image_generation(scheduler_output)
        return super().execute_model(scheduler_output, *args, **kwargs)   # standard LLM execution


# ------------------------------------------------------------------ lazy binding onto vLLM
_CLASSES: dict = {}


def _vllm_bases():
    try:
        from vllm.v1.executor.abstract import UniProcExecutor
        from vllm.v1.worker.gpu_worker import Worker
    except Exception as e:  # pragma: no cover - depends on the environment
        raise ImportError("vLLM V1 (vllm.v1.worker.gpu_worker / vllm.v1.executor.abstract) is required for the "
                          "serving wiring") from e
    return Worker, UniProcExecutor


def make_vllm_classes() -> dict:
    """Build (once) the four vLLM-bound classes: QwenImageWorker, QwenImageUniProcExecutor,
    QwenImageGenWorker, QwenImageGenExecutor."""
    if _CLASSES:
        return _CLASSES
    Worker, UniProcExecutor = _vllm_bases()

    class QwenImageWorker(QwenImageWorkerMixin, Worker):
        def init_device(self) -> None:
            super().init_device()
            self._qwen_image_adapter = None

        def load_model(self) -> None:
            super().load_model()
            self._load_qwen_vl()
            self.get_qwen_image_adapter()

    class QwenImageGenWorker(QwenImageGenWorkerMixin, Worker):
        def load_model(self) -> None:
            super().load_model()
            self.verify_qwen_components()

    class QwenImageUniProcExecutor(UniProcExecutor):
        def _init_executor(self) -> None:
            inject_worker_cls(self.vllm_config, QWEN_IMAGE_WORKER_CLS)
            super()._init_executor()

    class QwenImageGenExecutor(UniProcExecutor):
        def _init_executor(self) -> None:
            inject_worker_cls(self.vllm_config, QWEN_IMAGE_GEN_WORKER_CLS)
            super()._init_executor()

        def get_worker_class(self):
            return QwenImageGenWorker

    for c in (QwenImageWorker, QwenImageGenWorker, QwenImageUniProcExecutor, QwenImageGenExecutor):
        c.__module__ = __name__
        c.__qualname__ = c.__name__
        _CLASSES[c.__name__] = c
    return _CLASSES


def __getattr__(name: str):
    """``vllm_omni_draft_b200.serving.QwenImageWorker`` etc. resolve on first access, so that vLLM can import
    the worker by its dotted name without this package importing vLLM eagerly."""
    if name in ("QwenImageWorker", "QwenImageGenWorker", "QwenImageUniProcExecutor", "QwenImageGenExecutor"):
        return make_vllm_classes()[name]
    raise AttributeError(name)
