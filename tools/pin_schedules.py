"""Pin vllm_omni_draft_b200/tuned_schedules.json to the GEMM source it was measured on.

The in-tree schedule file is only valid for the kernels it was timed with.  After a change to csrc/qi_gemm.cu or
csrc/qi_common.cuh: bump qi_version() (csrc/qi_api.cu), re-tune on a B200 (any bench.py / pipeline run writes the
measured schedules to ~/.cache/vllm_omni_draft_b200/tuned_schedules.json), then run

    python tools/pin_schedules.py [measured.json ...]

which merges the given files' entries of the CURRENT version into the in-tree file, drops entries of other
versions and records the source hash that tests/test_cpu.py checks."""
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "vllm_omni_draft_b200")
PIN_KEY = "__gemm_source_sha256__"
PINNED_SOURCES = ("csrc/qi_gemm.cu", "csrc/qi_common.cuh")


def gemm_source_sha256() -> str:
    h = hashlib.sha256()
    for f in PINNED_SOURCES:
        with open(os.path.join(PKG, f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()


def main(argv) -> None:
    sys.path.insert(0, ROOT)
    from vllm_omni_draft_b200 import _lib
    tag = f"/v{_lib.lib().qi_version()}/"
    path = os.path.join(PKG, "tuned_schedules.json")
    merged = {}
    for f in [path] + list(argv):
        with open(f) as fh:
            merged.update({k: v for k, v in json.load(fh).items() if tag in k})
    merged[PIN_KEY] = gemm_source_sha256()
    with open(path, "w") as fh:
        json.dump(merged, fh, indent=0, sort_keys=True)
    print(f"{path}: {len(merged) - 1} schedules at {tag.strip('/')}, pinned to {merged[PIN_KEY][:12]}")


if __name__ == "__main__":
    main(sys.argv[1:])
