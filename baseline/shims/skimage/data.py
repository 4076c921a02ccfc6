def coffee():
    raise RuntimeError("skimage shim: sample images are not available offline")
