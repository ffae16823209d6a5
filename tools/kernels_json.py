#!/usr/bin/env python
"""Print bench.py's `roofline` / `kernels` table alone (per-kernel CUDA-event timings + library bars)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from bench import load_peaks  # noqa: E402
from vllm_omni_draft_b200.roofline import kernel_rooflines  # noqa: E402

if __name__ == "__main__":
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    torch.cuda.set_device(0)
    burst, sustained, hbm, src = load_peaks()
    print(json.dumps(kernel_rooflines(burst, sustained, hbm, src, batch=batch), indent=1))
