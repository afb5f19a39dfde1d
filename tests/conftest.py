"""pytest configuration: `gpu` marker, repo on sys.path, artefacts built on demand."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # The oracle is test infrastructure: build it here if a fresh checkout lacks it.
    if not os.path.exists(os.path.join(ROOT, "oracle", "libtk_oracle.so")):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "libtk_oracle.so"], check=True)
    if not os.path.exists(os.path.join(ROOT, "trackiellm_b200", "libtk_b200.so")):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "trackiellm_b200", "csrc")], check=True)


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
