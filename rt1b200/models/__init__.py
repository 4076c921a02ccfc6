from .film_efficientnet import (EfficientNet, EfficientNetB3, EfficientNetEncoder, FilmConditioning,  # noqa: F401
                                MBConvBlock, SeModule, round_filters, round_repeats)
from .token_learner import TokenLearnerModule  # noqa: F401
from .tokenizers import RT1ActionTokenizer, RT1ImageTokenizer, batched_space_sampler, np_to_tensor  # noqa: F401
from .transformer import TF_MultiHeadAttention, Transformer, attention  # noqa: F401
from .transformer_network import TransformerNetwork  # noqa: F401
from .factory import build_rt1, default_spaces, RT1Config  # noqa: F401
