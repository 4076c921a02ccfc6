"""Test-harness stand-in: see README.md."""


def getDataPath():
    return ""
