"""PyTorch restatement of the Language-Table LAVA baseline networks (reference ``language_table/train/networks``, Flax)."""
