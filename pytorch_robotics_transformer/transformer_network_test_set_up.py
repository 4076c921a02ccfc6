"""Test fixtures for ``TransformerNetwork`` under the reference's import path
(``pytorch_robotics_transformer/transformer_network_test_set_up.py``): the named state spaces, matching constant
observations for training ([b, t, ...]) and inference ([b, ...]), and a ``unittest`` base class that defines the action
space, the network-state space and fake action batches.  Users who wrote tests against the reference's helper keep their
imports; the implementations are this repository's (``rt1b200.spaces``, ``rt1b200.models``)."""
import unittest
from collections import OrderedDict
from typing import Dict, List

import numpy as np
import torch

from rt1b200 import spaces

BATCH_SIZE = 2
TIME_SEQUENCE_LENGTH = 3
HEIGHT = 256
WIDTH = 320
NUM_IMAGE_TOKENS = 2


def space_names_list() -> List[str]:
    return ["default"]


def state_space_list() -> List[spaces.Dict]:
    return [spaces.Dict({
        "image": spaces.Box(low=0.0, high=1.0, shape=(3, HEIGHT, WIDTH), dtype=np.float32),
        "natural_language_embedding": spaces.Box(low=-np.inf, high=np.inf, shape=[512], dtype=np.float32),
    })]


def observations_list(training: bool = True) -> List[Dict[str, torch.Tensor]]:
    lead = [BATCH_SIZE, TIME_SEQUENCE_LENGTH] if training else [1]          # the reference infers one environment at a time
    return [{"image": torch.full(lead + [3, HEIGHT, WIDTH], 0.5), "natural_language_embedding": torch.full(lead + [512], 1.0)}]


NAME_TO_STATE_SPACES = dict(zip(space_names_list(), state_space_list()))
NAME_TO_OBSERVATIONS = dict(zip(space_names_list(), observations_list()))
NAME_TO_INF_OBSERVATIONS = dict(zip(space_names_list(), observations_list(False)))


try:                                    # the reference's tests decorate their cases with absl's parameterized
    from absl.testing import parameterized as _parameterized
    _Base = (_parameterized.TestCase, unittest.TestCase)
except ImportError:                     # pragma: no cover - absl is optional here
    _Base = (unittest.TestCase,)


class TransformerNetworkTestUtils(*_Base):
    """``setUp`` defines what the reference's fixture defines: ``_state_space`` (observation space), ``_action_space`` (ordered:
    terminate_episode, world_vector, rotation_delta, gripper_closedness_action), ``_policy_info_space``, constant
    ``_train_observation`` / ``_train_action`` of shape [b, t, ...] and ``_inference_observation`` / ``_inference_action`` of
    shape [b, ...]."""

    def _define_spaces(self, train_batch_size=BATCH_SIZE, inference_batch_size=1, time_sequence_length=TIME_SEQUENCE_LENGTH,
                       inference_sequence_length=TIME_SEQUENCE_LENGTH, token_embedding_size=512, image_width=WIDTH,
                       image_height=HEIGHT):
        self.train_batch_size, self.inference_batch_size = train_batch_size, inference_batch_size
        self.time_sequence_length, self.inference_sequence_length = time_sequence_length, inference_sequence_length
        self.token_embedding_size = token_embedding_size
        box = lambda lo, hi, shape: spaces.Box(low=lo, high=hi, shape=shape, dtype=np.float32)      # noqa: E731
        # ordered: the action tokenizer walks the sub-actions in this order
        self._action_space = spaces.Dict(OrderedDict([
            ("terminate_episode", spaces.Discrete(2)), ("world_vector", box(-1.0, 1.0, (3,))),
            ("rotation_delta", box(-np.pi / 2, np.pi / 2, (3,))), ("gripper_closedness_action", box(-1.0, 1.0, (1,)))]))
        self._state_space = spaces.Dict({"image": box(0.0, 1.0, (3, image_height, image_width)),
                                         "natural_language_embedding": box(-np.inf, np.inf, [token_embedding_size])})
        self._policy_info_space = {"return": box(0.0, 1.0, ()), "discounted_return": box(0.0, 1.0, ())}

        def observation(lead, image_value):
            return {"image": torch.full(lead + [3, image_height, image_width], image_value),
                    "natural_language_embedding": torch.full(lead + [token_embedding_size], 1.0)}

        def action(lead):
            return {"world_vector": torch.full(lead + [3], 0.5), "rotation_delta": torch.full(lead + [3], 0.5),
                    "terminate_episode": torch.full(lead, 1), "gripper_closedness_action": torch.full(lead + [1], 0.5)}

        self._inference_observation = observation([inference_batch_size], 1.0)
        self._train_observation = observation([train_batch_size, time_sequence_length], 0.5)
        self._inference_action = action([inference_batch_size])
        self._train_action = action([train_batch_size, time_sequence_length])

    def setUp(self):
        self._define_spaces()
        super().setUp()
