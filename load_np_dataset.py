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


def main(argv=None):
    """``python load_np_dataset.py --data_dir <root>/train`` — the reference's dataset check (its ``__main__``,
    ``load_np_dataset.py:150-181``): build the windowed dataset over the given episode ids, pull one shuffled batch through
    the loader, print the shapes a training step receives and save one frame of it as a PNG to look at."""
    import argparse
    import glob
    import os

    import numpy as np
    from torch.utils.data import DataLoader

    ap = argparse.ArgumentParser(description=main.__doc__)
    ap.add_argument("--data_dir", required=True, help="directory holding episode_<n>.npy files (one split of the converter's output)")
    ap.add_argument("--num_episodes", type=int, default=0, help="use the first N episodes (0 = every file found)")
    ap.add_argument("--window_length", type=int, default=6)
    ap.add_argument("--batch_size", type=int, default=8)
    ap.add_argument("--random_crop_factor", type=float, default=0.95)
    ap.add_argument("--width", type=int, default=456)
    ap.add_argument("--height", type=int, default=256)
    ap.add_argument("--save_image", default="./imgs.png", help="where to write one frame of the batch ('' = nowhere)")
    a = ap.parse_args(argv)
    found = sorted(int(os.path.basename(f)[len("episode_"):-4]) for f in glob.glob(os.path.join(a.data_dir, "episode_*.npy")))
    if not found:
        raise SystemExit(f"no episode_<n>.npy under {a.data_dir}")
    ids = found[:a.num_episodes] if a.num_episodes else found
    debug(a.data_dir, ids[0], 0)
    ds = EmbodiedIntelligenceDataset(a.data_dir, ids, a.window_length,
                                     transform=DecodeAndRandomResizedCrop(a.random_crop_factor, (a.width, a.height)))
    batch = next(iter(DataLoader(ds, batch_size=min(a.batch_size, len(ds)), shuffle=True, collate_fn=collate_fn)))
    obs, label = batch["train_observation"], batch["action_label"]
    print(f"{len(ids)} episodes, {len(ds)} windows of {a.window_length} frames")
    print("image", tuple(obs["image"].shape), obs["image"].dtype)                         # (batch, window, 3, height, width)
    print("natural_language_embedding", tuple(obs["natural_language_embedding"].shape))   # (batch, window, 512)
    print("terminate_episode", tuple(label["terminate_episode"].shape), "action", tuple(label["action"].shape))
    if a.save_image:
        from rt1b200.utils.video import write_apng
        frame = (obs["image"][-1, -1].permute(1, 2, 0).clamp(0, 1) * 255).round().byte().numpy()
        print("wrote", write_apng(a.save_image, frame[None]))
    return batch


if __name__ == "__main__":
    main()
