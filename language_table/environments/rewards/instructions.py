"""All instructions of the six goal-conditioned task families (reference ``rewards/instructions.py:28-43``)."""
from language_table.environments.rewards import (block2absolutelocation, block2block, block2block_relative_location,
                                                 block2relativelocation, point2block, separate_blocks)

CLIP_VOCAB_SIZE = 49408


def generate_all_instructions(block_mode):
    out = []
    for family in (block2block, point2block, block2relativelocation, block2absolutelocation,
                   block2block_relative_location, separate_blocks):
        out += family.generate_all_instructions(block_mode)
    return out


def vocab_size(block_mode):
    return len({w for inst in generate_all_instructions(block_mode) for w in inst.split(" ")})
