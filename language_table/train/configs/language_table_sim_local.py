"""Local BC training of LAVA on simulated Language-Table data (reference
``language_table/train/configs/language_table_sim_local.py:28-81``), conv/max-pool image encoder."""
from language_table.examples.dataset_example import dataset_paths  # noqa: F401  (the upstream buckets, for reference)
from language_table.train.configs import ConfigDict


def get_config():
    config = ConfigDict()
    config.binary = "language_table/train/main"
    config.sequence_length = 4

    config.model_name = "sequence_lav_mse"
    config.model = ConfigDict()
    config.model.dense_resnet_width = 1024
    config.model.dense_resnet_num_blocks = 2
    config.model.lava_sequence_length = config.sequence_length
    config.model.lava_num_layers = 4
    config.model.lava_temporal_transformer_num_layers = 2
    config.model.lava_d_model = 128
    config.model.lava_num_heads = 2
    config.model.lava_pyramid_fuse_layers = (2, 3, 4)
    config.model.lava_image_encoder = "conv_maxpool"
    # the reference runs a frozen CLIP text tower in-graph ("clip"); offline the stored sentence embedding is used
    config.model.lava_lang_encoder = "clip_in_obs"

    config.agent_name = "bc"
    config.agent = ConfigDict()
    config.agent.learning_rate = 1e-3
    # [(file, [(prefix in the file, prefix in the model)])]; the reference points this at a converted CLIP text tower
    # (``download_clip_flax_ckpt.py``) when ``lava_lang_encoder = "clip"``
    config.agent.pretrained_checkpoints = []
    config.agent.freeze_keys = ["text_encoder"]

    config.dataset_path = None               # directory with {train,val,test}/episode_N.npy; None → synthetic episodes
    config.data_target_width = 320
    config.data_target_height = 180
    config.image_photometric_distortions = True
    config.image_augment_crop = True
    config.random_crop_factor = 0.95
    config.data_normalization_num_samples = 32
    config.data_skip_normalize_keys = ["rgb", "instruction"]
    config.synthetic_data = False

    config.num_train_steps = 1_000_000
    config.per_device_batch_size = 4         # "4096 is used for 64 TPUs" upstream
    config.replay_capacity = 5_000
    config.num_steps_per_train_iter = 1
    config.log_loss_every_steps = 50
    config.checkpoint_every_steps = 50
    config.seed = 42
    config.trial = 0
    return config


def get_hyper(h):
    """Sweep definition for a launcher (reference ``:84-89``): one trial.  ``h`` is any object with ``product`` and
    ``sweep`` — ``language_table.train.configs.Hyper`` is the local one."""
    return h.product([h.sweep("config.trial", range(1))])
