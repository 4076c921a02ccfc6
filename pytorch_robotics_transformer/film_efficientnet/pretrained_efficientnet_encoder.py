from rt1b200.models.film_efficientnet import EfficientNetEncoder  # noqa: F401
