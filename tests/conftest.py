import os
import sys

import pytest
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200 box)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    has_gpu = torch.cuda.is_available()
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords and not has_gpu:
            item.add_marker(skip_gpu)
        if "multigpu" in item.keywords and (not has_gpu or torch.cuda.device_count() < 2):
            item.add_marker(pytest.mark.skip(reason="needs >= 2 GPUs"))


class _Reference:
    """Handle on the unmodified reference modules (baseline/_ref), imported with the offline shims."""

    def __init__(self):
        sys.path.insert(0, os.path.join(REPO, "baseline"))
        import refenv
        sys.path.pop(0)
        if not refenv.available() and not refenv.try_install():
            pytest.skip(refenv.why_unavailable())
        saved_path = list(sys.path)
        tops = ("pytorch_robotics_transformer", "distribute_train", "load_np_dataset", "language_table")   # names both trees define
        saved_modules = {n: m for n, m in sys.modules.items() if n.split(".")[0] in tops}
        try:
            refenv.activate()
            import pytorch_robotics_transformer.transformer_network as tn
            import pytorch_robotics_transformer.transformer as tr
            import pytorch_robotics_transformer.tokenizers.action_tokenizer as at
            import pytorch_robotics_transformer.tokenizers.token_learner as tl
            import pytorch_robotics_transformer.film_efficientnet.film_efficientnet_encoder as fe
            import pytorch_robotics_transformer.film_efficientnet.preprocessors as pp
            import gym.spaces as gs
            self.transformer_network, self.transformer, self.action_tokenizer = tn, tr, at
            self.token_learner, self.film_efficientnet_encoder, self.preprocessors, self.spaces = tl, fe, pp, gs
        finally:
            sys.path[:] = saved_path
            # activate() evicted this repository's same-named packages from sys.modules; put them back (and drop the
            # reference's entries — its module objects stay alive as attributes of this handle), otherwise a later
            # ``import language_table...`` would create a SECOND copy of our modules next to the ones already bound.
            for n in [n for n in sys.modules if n.split(".")[0] in tops]:
                del sys.modules[n]
            sys.modules.update(saved_modules)


@pytest.fixture(autouse=True)
def _reset_dropout_seed_epoch():
    """A captured training step registers a device word that is added to every dropout seed (rt1_set_seed_epoch); tests that
    rebuild dropout masks on the host expect it to be zero."""
    yield
    try:
        from rt1b200.ops import kernels
        for t in kernels._seed_epoch.values():
            t.zero_()
    except Exception:
        pass


@pytest.fixture(scope="session")
def reference():
    return _Reference()


@pytest.fixture()
def cuda_literal_to_cpu(monkeypatch):
    """The reference hard-codes ``.to('cuda')`` (transformer.py:91, action_tokenizer.py:117); on a CPU-only
    box redirect that literal to the CPU so it can serve as the oracle."""
    if torch.cuda.is_available():
        yield
        return
    orig = torch.Tensor.to

    def to(self, *a, **k):
        a = tuple("cpu" if (isinstance(x, str) and x.startswith("cuda")) else x for x in a)
        if isinstance(k.get("device"), str) and k["device"].startswith("cuda"):
            k["device"] = "cpu"
        return orig(self, *a, **k)

    monkeypatch.setattr(torch.Tensor, "to", to)
    yield
