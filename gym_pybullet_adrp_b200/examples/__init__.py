"""The reference's example scripts (gym_pybullet_adrp/examples) on the batched envs: same ``run(...)`` keywords and command
lines, plus ``num_envs`` -- every env flies the same scenario side by side, env 0 is the one logged."""
