import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_sessionstart(session):
    """The built libraries are git-ignored: on a fresh checkout build them once (what __graft_entry__.build() does),
    so that `pytest tests` does not depend on a build step having run first.  Nothing is built if they exist."""
    import shutil
    import subprocess
    have = (os.path.exists(os.path.join(ROOT, "langsplat_b200", "liblsr.so"))
            and any(f.endswith(".so") for f in os.listdir(os.path.join(ROOT, "oracle"))))
    nvcc = shutil.which("nvcc") or ("/usr/local/cuda/bin/nvcc" if os.path.exists("/usr/local/cuda/bin/nvcc") else None)
    if have or nvcc is None:
        return
    env = dict(os.environ)
    env.pop("CXX", None)
    env.pop("CC", None)
    for d in ("langsplat_b200/csrc", "oracle"):
        subprocess.call(["make", "-C", os.path.join(ROOT, d), "-j8", "-s"], env=env)


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
