"""CUDA-graph timing of each hot kernel alone at the cfg-2 shapes (quick iteration aid)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_omni_draft_b200.roofline import kernel_rooflines  # noqa: E402

print(json.dumps(kernel_rooflines(1623.3, 1400.4, 6541.5, "measured"), indent=1))
