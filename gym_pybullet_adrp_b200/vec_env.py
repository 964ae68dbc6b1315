"""Torch-native vectorised view of a batched env for SB3-style on-policy loops.

The reference trains with Stable-Baselines3 on ONE ``HoverAviary`` / wrapped ``MultiRaceAviary`` per process
(``gym.make`` + ``make_vec_env``); every ``VecEnv.step`` there is a Python loop over env copies.  Here the batch IS
the env: ``num_envs`` envs advance in one kernel launch and observations, rewards and done flags stay on the
device as torch tensors.  The class follows the ``VecEnv`` calling convention (``reset() -> obs``,
``step(actions) -> obs, rewards, dones, infos``, ``step_async`` / ``step_wait``, ``num_envs``,
``observation_space`` / ``action_space`` of ONE env) so that a PPO rollout loop written against SB3's ``VecEnv``
runs unchanged on tensors.  Episode ends follow SB3's ``VecEnv`` contract: a finished env is reset inside ``step`` (one
masked ``adrp_reset`` launch per step, mask on the device, no host synchronisation), the observation returned for it is
the FIRST one of its next episode (for the race env: of the re-randomised track), and the last observation of the
finished episode is in ``infos["terminal_observation"]`` (rows of envs that are not done are just this step's rows).
The env must be built with ``auto_reset=False``: the in-kernel auto-reset of ``adrp_step`` re-initialises the state AFTER
writing the terminal row and never shows the reset observation, which is right for throughput rollouts and wrong here.
Returned tensors are copies: they stay valid across later steps (``Handle.step`` itself returns the handle's persistent
output tensors, which the next step overwrites in place).
"""
from __future__ import annotations

import numpy as np


class TorchVecEnv:
    def __init__(self, env, flatten: bool = True):
        base = getattr(env, "unwrapped", env)
        if base.NUM_ENVS < 2:
            raise ValueError("TorchVecEnv needs a batched env (num_envs > 1)")
        if base._auto_reset:
            raise ValueError("TorchVecEnv resets finished envs itself (so that it can return their first observation): "
                             "construct the env with auto_reset=False")
        self.env, self.base = env, base
        self.num_envs = base.NUM_ENVS
        self.flatten = flatten
        self.observation_space, self.action_space = base.observation_space, base.action_space
        self._actions = None
        self.episode_returns = None
        self.episode_lengths = None

    # -- VecEnv calling convention
    def reset(self):
        obs, _ = self.env.reset()
        import torch
        self.episode_returns = torch.zeros(self.num_envs, dtype=obs.dtype, device=obs.device)
        self.episode_lengths = torch.zeros(self.num_envs, dtype=torch.int64, device=obs.device)
        return self._shape(obs).clone()

    def step_async(self, actions):
        self._actions = actions

    def step_wait(self):
        import torch
        a = self._actions
        if not torch.is_tensor(a):
            a = torch.as_tensor(np.asarray(a))
        a = a.reshape((self.num_envs,) + tuple(self.action_space.shape))
        obs, reward, terminated, truncated, _ = self.env.step(a)
        dones = terminated | truncated
        reward = reward.clone()
        self.episode_returns += reward
        self.episode_lengths += 1
        # SB3 keeps per-env info dicts; on the device the same facts are tensors
        infos = {"TimeLimit.truncated": truncated & ~terminated, "episode_return": self.episode_returns.clone(),
                 "episode_length": self.episode_lengths.clone(), "terminal_observation": self._shape(obs).clone()}
        self.episode_returns = torch.where(dones, torch.zeros_like(self.episode_returns), self.episode_returns)
        self.episode_lengths = torch.where(dones, torch.zeros_like(self.episode_lengths), self.episode_lengths)
        # Env.reset for the finished envs only: writes their reset observation into the handle's obs rows
        obs = self.base.handle.reset(env_mask=dones)
        return self._shape(obs).clone(), reward, dones, infos

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def close(self):
        self.env.close()

    def _shape(self, obs):
        return obs.reshape(self.num_envs, -1) if self.flatten else obs
