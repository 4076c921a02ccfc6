from rt1b200.models.tokenizers import batched_space_sampler, np_to_tensor  # noqa: F401
