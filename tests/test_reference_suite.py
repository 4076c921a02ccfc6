"""The reference's OWN unit tests, run against this repository's implementation.

``pytorch_robotics_transformer`` here is a package of re-exports of ``rt1b200.models``; the reference's test files import the
model through that path, so they can be pointed at our classes unchanged.  Each file is executed in a fresh interpreter whose
``sys.path`` has this repository first (the test asserts that the package that got imported is ours).  Runnable offline: the
tokenizer, TokenLearner, decoder and TransformerNetwork tests (the causality test drives ``_transformer_call`` directly), and
three Language-Table files — the instruction-count test, the save / restore test of the environment and the xArm kinematics
test.  Those three import TensorFlow only for ``tf.test.TestCase`` and (the arm test) PyBullet only to open a client that
this implementation ignores; ``tests/shims`` supplies exactly that much (see its README).  Not runnable here: the two
EfficientNet tests (they need the ImageNet weight blob, which upstream does not ship either), ``resnet_v1_test`` (Flax) and
``clip_tokenizer_test`` (TF-Text and the ``clip`` package)."""
import os
import subprocess
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ROOTS = (os.path.join(REPO, "baseline", "_ref"), "/root/reference")
FILES = ["pytorch_robotics_transformer/tokenizers/action_tokenizer_test.py",
         "pytorch_robotics_transformer/tokenizers/token_learner_test.py",
         "pytorch_robotics_transformer/tokenizers/image_tokenizer_test.py",
         "pytorch_robotics_transformer/transformer_test.py",
         "pytorch_robotics_transformer/transformer_network_test.py"]

LT_FILES = ["language_table/environments/rewards/instructions_test.py",
            "language_table/environments/language_table_test.py",
            "language_table/environments/utils/xarm_sim_robot_test.py"]

RUNNER = r"""
import sys, runpy, unittest
sys.argv = ['reference-test']
import PKG as pkg
assert pkg.__file__.startswith(REPO), pkg.__file__
mod = runpy.run_path(PATH, run_name='reference_test')
suite = unittest.TestSuite()
for v in mod.values():
    if isinstance(v, type) and issubclass(v, unittest.TestCase) and any(n.startswith('test') for n in dir(v)):
        suite.addTests(unittest.defaultTestLoader.loadTestsFromTestCase(v))
r = unittest.TextTestRunner(verbosity=0).run(suite)
print('RESULT', r.testsRun, len(r.failures), len(r.errors))
for _, e in (r.errors + r.failures)[:3]:
    print(e[-1500:])
"""


def _find(rel):
    for root in ROOTS:
        p = os.path.join(root, rel)
        if os.path.isfile(p):
            return p
    return None


@pytest.mark.parametrize("rel", FILES)
def test_reference_unit_tests_pass_on_this_implementation(rel, tmp_path):
    path = _find(rel)
    if path is None:
        pytest.skip("the reference's test files are not on this machine")
    try:
        import absl.testing.parameterized  # noqa: F401
    except ImportError:
        pytest.skip("absl is not installed")
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([REPO, os.path.join(REPO, "baseline", "shims")]), CUDA_VISIBLE_DEVICES="")
    _run(path, "pytorch_robotics_transformer", env, tmp_path)


def _run(path, pkg, env, tmp_path):
    code = RUNNER.replace("REPO", repr(REPO)).replace("PATH", repr(path)).replace("PKG", pkg)
    r = subprocess.run([sys.executable, "-c", code], cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=900)
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT")]
    assert line, (r.stdout[-1500:], r.stderr[-1500:])
    ran, failures, errors = (int(x) for x in line[0].split()[1:])
    assert ran > 0 and failures == 0 and errors == 0, (r.stdout[-3000:], r.stderr[-1500:])
    return ran


@pytest.mark.parametrize("rel", LT_FILES)
def test_reference_language_table_tests_pass_on_this_implementation(rel, tmp_path):
    path = _find(rel)
    if path is None:
        pytest.skip("the reference's test files are not on this machine")
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([REPO, os.path.join(REPO, "tests", "shims")]), CUDA_VISIBLE_DEVICES="")
    _run(path, "language_table", env, tmp_path)


def test_reference_environment_example_runs_on_this_implementation(tmp_path):
    """The reference's ``examples/environment_example.py`` — build the BLOCK_8 block-to-block environment, reset, five random
    actions from ``env.action_space``, save ``env.render()`` — executed unchanged against this package."""
    path = _find("language_table/examples/environment_example.py")
    if path is None:
        pytest.skip("the reference's files are not on this machine")
    try:
        import absl.app  # noqa: F401
    except ImportError:
        pytest.skip("absl is not installed")
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([REPO, os.path.join(REPO, "tests", "shims")]), CUDA_VISIBLE_DEVICES="")
    out = "/tmp/language_table_render.png"                                 # the path is fixed in the example
    if os.path.exists(out):
        os.remove(out)
    code = ("import sys, runpy; import language_table as p; assert p.__file__.startswith(%r); sys.argv = ['example']; "
            "runpy.run_path(%r, run_name='__main__')" % (REPO, path))
    r = subprocess.run([sys.executable, "-c", code], cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    data = open(out, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n" and len(data) > 10_000       # the instruction band + the camera frame
