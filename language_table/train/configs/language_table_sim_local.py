"""Local BC training of LAVA on simulated Language-Table data (reference
``language_table/train/configs/language_table_sim_local.py:28-81``), conv/max-pool image encoder."""
import copy

from language_table.examples.dataset_example import dataset_paths  # noqa: F401  (the upstream buckets, for reference)
from language_table.train.configs import ConfigDict


SEQUENCE_LENGTH = 4

# the model: LAVA over `SEQUENCE_LENGTH` frames, conv / max-pool image pyramid, a dense-ResNet action head
MODEL = dict(dense_resnet_width=1024, dense_resnet_num_blocks=2, lava_sequence_length=SEQUENCE_LENGTH, lava_num_layers=4,
             lava_temporal_transformer_num_layers=2, lava_d_model=128, lava_num_heads=2, lava_pyramid_fuse_layers=(2, 3, 4),
             lava_image_encoder="conv_maxpool",
             # the reference runs a frozen CLIP text tower in-graph ("clip"); offline the stored sentence embedding is used
             lava_lang_encoder="clip_in_obs")

# the agent: behaviour cloning; `pretrained_checkpoints` = [(file, [(prefix in the file, prefix in the model)])] — the
# reference points it at a converted CLIP text tower (``download_clip_flax_ckpt.py``) when lava_lang_encoder = "clip"
AGENT = dict(learning_rate=1e-3, pretrained_checkpoints=[], freeze_keys=["text_encoder"])

# data: directory with {train,val,test}/episode_N.npy (None → synthetic episodes), 320 × 180 frames, a 0.95 crop window and
# photometric distortions in training, statistics over 32 batches, images and instructions not normalised
DATA = dict(dataset_path=None, clip_bpe_path=None,          # CLIP's bpe_simple_vocab_16e6.txt.gz, for lava_lang_encoder = "clip"
            data_target_width=320, data_target_height=180, image_photometric_distortions=True,
            image_augment_crop=True, random_crop_factor=0.95, data_normalization_num_samples=32,
            data_skip_normalize_keys=["rgb", "instruction"], synthetic_data=False)

# the run: "4096 is used for 64 TPUs" upstream for the per-device batch
RUN = dict(num_train_steps=1_000_000, per_device_batch_size=4, replay_capacity=5_000, num_steps_per_train_iter=1,
           log_loss_every_steps=50, checkpoint_every_steps=50, seed=42, trial=0)


def get_config():
    config = ConfigDict(binary="language_table/train/main", sequence_length=SEQUENCE_LENGTH, model_name="sequence_lav_mse",
                        model=ConfigDict(MODEL), agent_name="bc", agent=ConfigDict(copy.deepcopy(AGENT)))
    config.update(copy.deepcopy(DATA))
    config.update(RUN)
    return config


def get_hyper(h):
    """Sweep definition for a launcher (reference ``:84-89``): one trial.  ``h`` is any object with ``product`` and
    ``sweep`` — ``language_table.train.configs.Hyper`` is the local one."""
    return h.product([h.sweep("config.trial", range(1))])
