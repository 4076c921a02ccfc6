"""CLIP text-tower checkpoint for the LAVA baseline's ``lang_encoder="clip"``.

The reference's script of this name downloads scenic's CLIP model variables and re-saves them as a clu checkpoint
(``language_table/train/download_clip_flax_ckpt.py``).  There is no network here and no scenic / clu; what can be done
offline is the second half: take a CLIP ``state_dict`` the user already has on disk (the public PyTorch checkpoints: a
``.pt`` file holding either the state dict or a model object with ``state_dict()``), keep the text tower, verify every
shape against :class:`language_table.train.networks.clip_text.ClipTextEncoder` and write it where the trainer looks.

    python -m language_table.train.download_clip_flax_ckpt --source /path/to/clip.pt --checkpoint_directory DIR
"""
import argparse
import os

import torch

from language_table.train.networks.clip_text import ClipTextEncoder, load_clip_text_weights


def convert(source: str, checkpoint_directory: str) -> str:
    obj = torch.load(source, map_location="cpu", weights_only=False)
    state = obj.state_dict() if hasattr(obj, "state_dict") else obj
    if "state_dict" in state and isinstance(state["state_dict"], dict):
        state = state["state_dict"]
    width = state["token_embedding.weight"].shape[1]
    layers = 1 + max(int(k.split(".")[2]) for k in state if k.startswith("transformer.resblocks."))
    tower = ClipTextEncoder(vocab_size=state["token_embedding.weight"].shape[0], features=width, num_heads=max(1, width // 64),
                            num_layers=layers, out_features=state["text_projection"].shape[1],
                            context_length=state["positional_embedding"].shape[0])
    load_clip_text_weights(tower, state)
    os.makedirs(checkpoint_directory, exist_ok=True)
    out = os.path.join(checkpoint_directory, "clip_text_tower.pt")
    torch.save(tower.state_dict(), out)
    return out


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--source", required=True, help="a CLIP checkpoint on disk (downloading is not possible offline)")
    ap.add_argument("--checkpoint_directory", required=True)
    a = ap.parse_args(argv)
    print("wrote", convert(a.source, a.checkpoint_directory))


if __name__ == "__main__":
    main()
