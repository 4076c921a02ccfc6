import unittest

import numpy as np


class TestCase(unittest.TestCase):
    def assertLen(self, container, expected_len, msg=None):
        self.assertEqual(len(container), expected_len, msg)

    def assertAllClose(self, a, b, rtol=1e-6, atol=1e-6, msg=None):
        np.testing.assert_allclose(np.asarray(a), np.asarray(b), rtol=rtol, atol=atol, err_msg=msg or "")

    def assertAllEqual(self, a, b, msg=None):
        np.testing.assert_array_equal(np.asarray(a), np.asarray(b), err_msg=msg or "")

    def assertEmpty(self, container, msg=None):
        self.assertEqual(len(container), 0, msg)


def main(*args, **kwargs):
    unittest.main(*args, **kwargs)
