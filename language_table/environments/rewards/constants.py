"""Constants shared by the board rewards (reference ``rewards/constants.py``)."""
TARGET_BLOCK_DISTANCE = 0.05
