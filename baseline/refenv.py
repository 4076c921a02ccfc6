"""Locate and activate the UNMODIFIED reference (``baseline/_ref``, installed from /root/reference with pip
``--target``; see DESIGN.md §Reference arm) plus the import shims it needs offline (``baseline/shims``:
``gym.spaces``, ``skimage``, ``matplotlib``, ``lightning``).

Nothing from ``rt1b200`` is imported by the reference arm except the space descriptors behind the ``gym``
shim (pure numpy).  ``activate()`` puts both directories at the FRONT of ``sys.path`` and removes the
repository root, so ``pytorch_robotics_transformer`` / ``distribute_train`` / ``language_table`` resolve to
the reference's files and not to this repository's same-named compatibility modules.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
REF = os.path.join(HERE, "_ref")
SHIMS = os.path.join(HERE, "shims")
_NOTOP = os.path.join(REF, "pytorch_robotics_transformer", "film_efficientnet", "efficientnet_checkpoints",
                      "efficientnetb3_notop.pth")


def available() -> bool:
    return os.path.isfile(os.path.join(REF, "distribute_train.py"))


def why_unavailable() -> str:
    return f"{REF} missing (pip --target install of /root/reference was not done on this box)"


def try_install() -> bool:
    """Re-do the offline install if /root/reference is mounted (it is on the build box, not on GPU boxes)."""
    src = "/root/reference"
    if available() or not os.path.isdir(src):
        return available()
    tmp = "/tmp/rt1_refcopy"
    subprocess.run(["rm", "-rf", tmp], check=False)
    subprocess.run(["cp", "-r", src, tmp], check=True)
    with open(os.path.join(tmp, "setup.py"), "w") as f:
        f.write(
            "from setuptools import setup, find_namespace_packages\n"
            "setup(name='pytorch-rt1-reference', version='0.0.0',\n"
            "      packages=find_namespace_packages(include=['pytorch_robotics_transformer*','language_table*']),\n"
            "      py_modules=['distribute_train','load_np_dataset','rlds_np_convert'],\n"
            "      include_package_data=True, package_data={'': ['*.json','*.urdf','*.obj','*.md']})\n")
    r = subprocess.run([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps",
                        "--find-links", "/opt/wheelhouse", "--target", REF, tmp], capture_output=True, text=True)
    return r.returncode == 0 and available()


def ensure_random_efficientnet_blob() -> str:
    """The reference's image tokenizer unconditionally loads ``efficientnetb3_notop.pth`` (ImageNet weights the
    upstream repo does not ship, ``.MISSING_LARGE_BLOBS``).  Benchmarks use random-init weights, so write a
    random-init torchvision EfficientNet-B3 feature state dict there: ``load_official_pytorch_param`` matches
    tensors by ORDER, which torchvision's layout satisfies.  This adds a data file beside the reference's
    sources; no reference source file is touched."""
    if not os.path.exists(_NOTOP):
        import torch
        import torchvision
        os.makedirs(os.path.dirname(_NOTOP), exist_ok=True)
        torch.manual_seed(0)
        sd = torchvision.models.efficientnet_b3(weights=None).features.state_dict()
        torch.save(sd, _NOTOP)
    return _NOTOP


def activate() -> None:
    if not available():
        raise RuntimeError(why_unavailable())
    ensure_random_efficientnet_blob()
    drop = {REPO, "", "."}
    sys.path[:] = [p for p in sys.path if os.path.abspath(p or ".") not in drop and p not in drop]
    sys.path.insert(0, REF)
    sys.path.insert(0, SHIMS)
    for name in list(sys.modules):
        if name.split(".")[0] in ("pytorch_robotics_transformer", "distribute_train", "load_np_dataset",
                                  "language_table", "gym", "lightning"):
            del sys.modules[name]
