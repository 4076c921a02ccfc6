"""Import-path compatibility with the reference's ``load_np_dataset.py``."""
from rt1b200.data.dataset import DecodeAndRandomResizedCrop, EmbodiedIntelligenceDataset, collate_fn  # noqa: F401


def debug(data_dir, episode_id, step_id):
    """Print what one step of one episode file holds (the reference's ``debug``, ``load_np_dataset.py:118-127``)."""
    import os

    import numpy as np
    step = np.load(os.path.join(data_dir, f"episode_{episode_id}.npy"), allow_pickle=True)[step_id]
    print(f"action: {step['action']}")
    print(f"is_first: {step['is_first']}")
    print(f"is_terminal: {step['is_terminal']}")
    print(f"instruction: {np.asarray(step['instruction']).shape}")
    print(f"rgb: {np.asarray(step['rgb']).shape}")
