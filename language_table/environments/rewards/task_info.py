"""Task descriptions returned by ``reward.reset`` (reference ``rewards/task_info.py:24-104``)."""
from dataclasses import dataclass
from typing import Any, List

FAILURE = "failure"


@dataclass
class Block2BlockTaskInfo:
    instruction: str
    block1: str
    block2: str


@dataclass
class Point2BlockTaskInfo:
    instruction: str
    block_target: str


@dataclass
class Block2LocationTaskInfo:
    instruction: str
    block: str
    location: str
    target_translation: Any


@dataclass
class Block2LineTaskInfo:
    """Push a block onto a line of the board (declared by the reference, ``task_info.py:42``; no reward uses it there either)."""
    instruction: str
    block: str
    target_translation: Any


@dataclass
class Block2PoleTaskInfo:
    """Push a block to one of the coloured poles (``task_info.py:50``; the poles are listed in ``blocks.POLE_URDF_PATHS``)."""
    instruction: str
    block1: str
    goal: str


@dataclass
class Block2RelativeLocationTaskInfo:
    instruction: str
    block: str
    location: str
    target_translation: Any


@dataclass
class Block2BlockRelativeLocationTaskInfo:
    instruction: str
    block: str
    target_block: str
    direction: str
    target_translation: Any


@dataclass
class SeparateBlocksTaskInfo:
    instruction: str
    block: str
    avoid_blocks: List[str]
    target_translation: Any


@dataclass
class PlayTaskInfo:
    instruction: str


ALL_TASKS = [Block2BlockTaskInfo, Point2BlockTaskInfo, Block2LocationTaskInfo, Block2RelativeLocationTaskInfo,
             Block2BlockRelativeLocationTaskInfo, SeparateBlocksTaskInfo, PlayTaskInfo]
