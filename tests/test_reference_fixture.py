"""The reference's TransformerNetwork test fixture, under its import path, drives this repository's network."""
import torch

from pytorch_robotics_transformer import transformer_network_test_set_up as fx
from pytorch_robotics_transformer.transformer_network import TransformerNetwork


class _Case(fx.TransformerNetworkTestUtils):
    def runTest(self):
        pass


def test_fixture_builds_a_network_that_trains_and_infers():
    case = _Case()
    case._define_spaces(image_width=96, image_height=64, time_sequence_length=2)
    assert list(case._action_space.keys())[0] == "terminate_episode"
    assert case._train_observation["image"].shape == (2, 2, 3, 64, 96) and case._inference_action["world_vector"].shape == (1, 3)
    net = TransformerNetwork(input_tensor_space=case._state_space, output_tensor_space=case._action_space, vocab_size=256,
                             token_embedding_size=512, num_layers=1, layer_size=16, num_heads=2, feed_forward_size=32,
                             dropout_rate=0.0, time_sequence_length=2, crop_size=60, use_token_learner=True)
    assert net._tokens_per_action == 8
    state = {"context_image_tokens": torch.zeros(2, 2, 8, 512), "action_tokens": torch.zeros(2, 2, 8, dtype=torch.long),
             "seq_idx": torch.zeros(2, dtype=torch.long)}
    net.set_actions(case._train_action)
    net.train()
    net(case._train_observation, state)
    loss = net.get_actor_loss().mean()
    loss.backward()
    assert torch.isfinite(loss)
    assert fx.NAME_TO_STATE_SPACES["default"]["image"].shape == (3, fx.HEIGHT, fx.WIDTH)
    assert fx.NAME_TO_INF_OBSERVATIONS["default"]["image"].shape == (1, 3, fx.HEIGHT, fx.WIDTH)
