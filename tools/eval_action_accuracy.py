"""Open-loop check of a checkpoint on held-out episodes: how well does the policy predict the demonstrated actions?

    python tools/eval_action_accuracy.py --ckpt ckpts/exp/last.ckpt --data_dir data/val [--height 256 --width 456]

Reports, over every window of the split (teacher-forced, evaluation mode, centre crop): the cross-entropy per action token, token
accuracy (exact and within ±2 bins), the mean absolute error of the de-tokenised action in metres — and the same figures for two
baselines computed from the same labels: always the most frequent token ("mode"), and the empirical token distribution
("marginal", the cross-entropy a policy that ignores its input converges to).  A model that sits at the marginal line has not
learnt to look at the image yet, whatever its training loss curve looks like (CPU-scale walk-through).
"""
import argparse
import glob
import json
import os
import sys
from argparse import Namespace

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--ckpt", required=True)
    ap.add_argument("--data_dir", required=True, help="directory with episode_<n>.npy (one split)")
    ap.add_argument("--height", type=int, default=256)
    ap.add_argument("--width", type=int, default=456)
    ap.add_argument("--seq_len", type=int, default=6)
    ap.add_argument("--batch_size", type=int, default=8)
    ap.add_argument("--max_batches", type=int, default=0)
    ap.add_argument("--device", default=None)
    a = ap.parse_args(argv)
    from torch.utils.data import DataLoader

    from load_np_dataset import DecodeAndRandomResizedCrop, EmbodiedIntelligenceDataset, collate_fn
    from rt1b200.train.trainer import load_policy_from_checkpoint
    device = a.device or ("cuda" if torch.cuda.is_available() else "cpu")
    module = load_policy_from_checkpoint(a.ckpt, Namespace(height=a.height, width=a.width, seq_len=a.seq_len), device=device,
                                         dtype=torch.bfloat16 if device.startswith("cuda") else None)
    model = module.model
    model._eval_center_crop = True
    ids = sorted(int(os.path.basename(f)[len("episode_"):-4]) for f in glob.glob(os.path.join(a.data_dir, "episode_*.npy")))
    ds = EmbodiedIntelligenceDataset(a.data_dir, ids, a.seq_len, transform=DecodeAndRandomResizedCrop(None, (a.width, a.height)))
    loader = DataLoader(ds, batch_size=a.batch_size, shuffle=False, collate_fn=collate_fn)
    labels, preds, ce_sum, n_tok = [], [], 0.0, 0
    with torch.no_grad():
        for i, batch in enumerate(loader):
            if a.max_batches and i >= a.max_batches:
                break
            obs = {k: v.to(device) for k, v in batch["train_observation"].items()}
            lab = {k: v.to(device) for k, v in batch["action_label"].items()}
            if device.startswith("cuda"):
                obs["natural_language_embedding"] = obs["natural_language_embedding"].to(torch.bfloat16)
            model.set_actions(lab)
            model(obs, None)
            aux = model.get_aux_info()
            y, p = aux["action_labels"].reshape(-1, aux["action_labels"].shape[-1]), aux["action_predictions"].reshape(-1, aux["action_labels"].shape[-1])
            b, t = lab["action"].shape[:2]
            ce_sum += float(model.get_actor_loss().float().sum()) * (b * t * model._single_time_step_num_tokens)   # undo /(b t 11), mean over A kept
            n_tok += y.shape[0]
            labels.append(y.cpu().numpy())
            preds.append(p.cpu().numpy())
    y, p = np.concatenate(labels), np.concatenate(preds)                   # [windows * T, A]; column 0 = terminate, 1.. = action dims
    vocab = model._vocab_size
    lo, hi = -0.1, 0.1
    space = getattr(model, "_action_space", None) or getattr(model, "_output_tensor_space", None)
    try:
        box = space["action"]
        lo, hi = float(np.min(box.low)), float(np.max(box.high))
    except Exception:
        pass
    to_m = lambda tok: tok / (vocab - 1) * (hi - lo) + lo                 # noqa: E731
    out = {"windows": len(ds), "steps": int(y.shape[0]), "action_range_m": [lo, hi],
           "model": {"ce_per_token": ce_sum / max(n_tok, 1)}}
    act = slice(1, y.shape[1])
    out["model"].update({"token_accuracy": float((p[:, act] == y[:, act]).mean()),
                         "token_accuracy_within_2": float((np.abs(p[:, act] - y[:, act]) <= 2).mean()),
                         "action_mae_m": float(np.abs(to_m(p[:, act]) - to_m(y[:, act])).mean()),
                         "terminate_accuracy": float((p[:, 0] == y[:, 0]).mean())})
    ce_marg, mode_acc, mode_mae, mode_w2 = [], [], [], []
    for j in range(y.shape[1]):
        counts = np.bincount(y[:, j], minlength=vocab).astype(np.float64)
        prob = counts / counts.sum()
        ce_marg.append(float(-(prob[prob > 0] * np.log(prob[prob > 0])).sum()))
        if j >= 1:
            mode = int(counts.argmax())
            mode_acc.append(float((y[:, j] == mode).mean()))
            mode_w2.append(float((np.abs(y[:, j] - mode) <= 2).mean()))
            mode_mae.append(float(np.abs(to_m(y[:, j]) - to_m(mode)).mean()))
    out["marginal"] = {"ce_per_token": float(np.mean(ce_marg))}
    out["mode"] = {"token_accuracy": float(np.mean(mode_acc)), "token_accuracy_within_2": float(np.mean(mode_w2)),
                   "action_mae_m": float(np.mean(mode_mae))}
    print(json.dumps(out, indent=1))
    return out


if __name__ == "__main__":
    main()
