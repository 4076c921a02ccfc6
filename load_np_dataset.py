"""Import-path compatibility with the reference's ``load_np_dataset.py``."""
from rt1b200.data.dataset import DecodeAndRandomResizedCrop, EmbodiedIntelligenceDataset, collate_fn  # noqa: F401
