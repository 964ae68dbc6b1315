"""The closed loop all three flight examples share (examples/pid.py:127-167, downwash.py:79-113, pid_velocity.py:113-146):
step the env, compute the next action from the observation, log env 0."""
from __future__ import annotations

import time

import numpy as np

from ..utils.logger import Logger
from ..utils.utils import sync


def refuse_graphics(gui, record_video, plot):
    if gui or record_video or plot:
        raise NotImplementedError("gui / record_video / plot need a graphics stack: unavailable on this box "
                                  "(pass --gui False --record_video False --plot False)")


def fly(env, policy, control_row, steps, *, output_folder, colab=False, csv_name="", gui=False, verbose=True):
    """``policy(i, obs) -> action`` with ``obs`` / ``action`` shaped like ``env.step``'s; ``control_row(i, j)`` = the
    12-float reference row the Logger stores for drone j at control step i.  Returns (logger, last observation)."""
    M = env.NUM_DRONES
    logger = Logger(logging_freq_hz=env.CTRL_FREQ, num_drones=M, output_folder=output_folder, colab=colab)
    batched = env.NUM_ENVS > 1
    action = np.zeros((env.NUM_ENVS, M, env.action_space.shape[-1])) if batched else np.zeros(env.action_space.shape)
    env.reset()
    start = time.time()
    obs = None
    for i in range(steps):
        obs, _, _, _, _ = env.step(action)
        obs_np = obs.cpu().numpy() if batched else obs
        action = policy(i, obs_np)
        first = obs_np[0] if batched else obs_np
        for j in range(M):
            logger.log(drone=j, timestamp=i / env.CTRL_FREQ, state=first[j], control=control_row(i, j))
        if verbose and i % env.CTRL_FREQ == 0:
            env.render()
        if gui:
            sync(i, start, env.CTRL_TIMESTEP)
    logger.save()
    if csv_name:
        logger.save_as_csv(csv_name)
    return logger, obs

