"""Test-harness stand-in: ``plt.imsave(path, uint8 image)`` through this repository's PNG writer — the one pyplot call the
reference's ``examples/environment_example.py`` makes."""
import numpy as np


def imsave(path, image, **kwargs):
    from rt1b200.utils.video import write_apng
    image = np.asarray(image)
    if image.dtype != np.uint8:
        image = (np.clip(image, 0.0, 1.0) * 255).round().astype(np.uint8)
    return write_apng(path, image[None])
