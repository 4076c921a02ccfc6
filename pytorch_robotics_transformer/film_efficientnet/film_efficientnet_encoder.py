from rt1b200.models.film_efficientnet import (CONV_KERNEL_INITIALIZER, DEFAULT_BLOCKS_ARGS, DENSE_KERNEL_INITIALIZER, EfficientNet, EfficientNetB3, ILSVRCPredictor, MBConvBlock, SeModule,  # noqa: F401
                                               load_official_pytorch_param, maybe_restore_with_film, round_filters, round_repeats)
