"""Create a board, run the scripted oracle for one episode, save the frames (reference ``examples/environment_example.py``)."""
import numpy as np

from language_table.environments import blocks, language_table
from language_table.environments.oracles import push_oracle_rrt_slowdown
from language_table.environments.rewards import block2block


def main(out="/tmp/language_table_example.npy", steps=80):
    env = language_table.LanguageTable(block_mode=blocks.LanguageTableBlockVariants.BLOCK_8,
                                       reward_factory=block2block.BlockToBlockReward, seed=0)
    obs = env.reset()
    oracle = push_oracle_rrt_slowdown.ObstacleOrientedPushOracleBoard2dRRT(env)
    frames = [obs["rgb"]]
    for _ in range(steps):
        obs, reward, done, _ = env.step(oracle.action())
        frames.append(obs["rgb"])
        if done:
            break
    np.save(out, np.stack(frames))
    print(f"instruction: {env.instruction_str!r}; success={env.succeeded}; {len(frames)} frames → {out}")


if __name__ == "__main__":
    main()
