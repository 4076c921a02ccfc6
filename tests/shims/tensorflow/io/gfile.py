"""Test-harness stand-in: ``tf.io.gfile`` over the local file system (the reference's state-file helpers open files through it)."""
import os


def GFile(name, mode="r"):
    return open(name, mode)


def makedirs(path):
    os.makedirs(path, exist_ok=True)


def exists(path):
    return os.path.exists(path)
