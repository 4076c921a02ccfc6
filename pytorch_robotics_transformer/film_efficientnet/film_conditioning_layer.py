from rt1b200.models.film_efficientnet import FilmConditioning  # noqa: F401
