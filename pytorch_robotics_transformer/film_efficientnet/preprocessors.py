from rt1b200.models.preprocessors import convert_dtype_and_crop_images  # noqa: F401
