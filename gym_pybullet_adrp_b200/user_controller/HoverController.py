"""``HoverController`` (user_controller/HoverController.py:8-27): a fixed setpoint per drone."""
import numpy as np

from gym_pybullet_adrp_b200.user_controller.BaseController import BaseController   # absolute: load_controller imports this file by path


class HoverController(BaseController):
    def predict(self, obs, reward=None, done=None, info=None, ep_time=None):
        return np.array([1.5, 0.5, 1 + self.drone_id, 0])

    @staticmethod
    def predict_batch(obs, ep_time=None):
        """The same setpoints for every env of the batch, as a ``(num_envs, NUM_DRONES, 4)`` tensor."""
        import torch
        N, M = obs.shape[0], obs.shape[1]
        a = torch.zeros((N, M, 4), dtype=obs.dtype, device=obs.device)
        a[..., 0], a[..., 1] = 1.5, 0.5
        a[..., 2] = 1 + torch.arange(M, dtype=obs.dtype, device=obs.device)
        return a
