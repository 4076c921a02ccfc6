"""As ``language_table_sim_local`` with the frozen-ResNet image pyramid (reference
``language_table/train/configs/language_table_resnet_sim_local.py``)."""
from language_table.train.configs import language_table_sim_local
from language_table.train.configs.language_table_sim_local import dataset_paths, get_hyper  # noqa: F401


def get_config():
    config = language_table_sim_local.get_config()
    config.model.lava_image_encoder = "resnet"
    config.agent.freeze_keys = ["text_encoder", "image_encoder.resnet"]
    return config
