"""``BaseController`` (user_controller/BaseController.py:10-117): the interface ``scripts/sim.py`` drives.

One instance per (episode, drone); ``predict(obs, ep_time=...)`` returns either a ``[x, y, z, yaw]`` setpoint or a
``(Command.FULLSTATE, args)`` tuple.  A controller that can act on the whole batch at once may additionally define
``predict_batch(obs, ep_time)`` on tensors of shape ``(num_envs, NUM_DRONES, obs_dim)`` -- ``scripts.sim.simulate``
then skips the per-drone Python loop.
"""
from abc import ABC, abstractmethod
from collections import deque

import numpy as np


class BaseController(ABC):
    def __init__(self, drone_id: int, initial_obs: np.ndarray = None, initial_info: dict = None, buffer_size: int = 100,
                 verbose: bool = False):
        self.drone_id, self.initial_obs, self.initial_info = drone_id, initial_obs, initial_info
        self.buffer_size, self.verbose = buffer_size, verbose
        self.reset()
        self.episode_reset()

    def reset(self):                                       # BaseController.py:50-56
        for name in ("action_buffer", "obs_buffer", "reward_buffer", "done_buffer", "info_buffer"):
            setattr(self, name, deque([], maxlen=self.buffer_size))

    def episode_reset(self):
        """Reset the controller's internal state and models if necessary."""

    def step_learn(self, action, obs, reward, done, info):  # :63-90
        self.action_buffer.append(action)
        self.obs_buffer.append(obs)
        self.reward_buffer.append(reward)
        self.done_buffer.append(done)
        self.info_buffer.append(info)

    def episode_learn(self):
        """Learning and controller updates called between episodes."""

    @abstractmethod
    def predict(self, obs: np.ndarray, reward: float = None, done: bool = None, info: dict = None, ep_time: float = None):
        """Predict the next action."""
