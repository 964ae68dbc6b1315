"""Batched evaluation driver with the interface of the reference's ``scripts/sim.py:18-112``.

``simulate(config, controller, n_runs, n_drones)`` returns the list of episode times like the reference does, but the
``n_runs`` episodes run side by side as ``n_runs`` envs of one device handle: every control step is ONE kernel launch
for all runs.  Controllers are the reference's ``BaseController`` subclasses (one instance per run and drone, called
on the host with that drone's observation row), or -- when the class defines ``predict_batch`` -- one call per step
on the whole ``(n_runs, n_drones, obs_dim)`` tensor with nothing leaving the device.

    python -m gym_pybullet_adrp_b200.scripts.sim --config config/level3.yaml \\
        --controller gym_pybullet_adrp_b200/user_controller/HoverController.py --n_runs 4096 --n_drones 2

``log_env`` dumps the flight of one run in the reference ``Logger``'s ``.npy`` layout (utils/logger.py).
"""
from __future__ import annotations

import argparse
import json
from typing import List

import numpy as np

from .. import _abi
from ..envs import MultiRaceAviary
from ..utils.enums import Command, Physics, RaceMode
from ..utils.logger import Logger
from ..utils.utils import load_config, load_controller


def _row(action, step_counter=0):
    """One controller output -> 16-float command row (MultiRaceAviary.py:190-205): a ``(Command, args)`` tuple of any
    Command, or an ndarray [x, y, z, yaw] = FULLSTATE(pos, 0, 0, yaw, 0, step_counter)."""
    if isinstance(action, tuple) and len(action) == 2 and isinstance(action[0], Command):
        return _abi.command_row(*action)
    a = np.asarray(action, float)
    return _abi.command_row(Command.FULLSTATE, [a[0:3], np.zeros(3), np.zeros(3), a[3], np.zeros(3), step_counter])


def simulate(config: str = "config/getting_started.yaml", controller: str | List[str] = None, n_runs: int = 10,
             n_drones: int = 2, gui: bool = False, *, racemode: RaceMode = RaceMode.COMPARE, physics: Physics = Physics.PYB,
             device: int = 0, precision: str = "fp32", seed: int = None, log_env: int = None,
             output_folder: str = "results", return_stats: bool = False):
    """Evaluate the drone controller(s) over ``n_runs`` episodes (scripts/sim.py:18-107).  Returns the episode times."""
    if gui:
        raise NotImplementedError("gui=True needs PyBullet's GUI: unavailable on this box")
    cfg = load_config(config) if isinstance(config, str) else config
    if controller is None:
        from ..user_controller import HoverController
        classes = [HoverController] * n_drones
    else:
        paths = [controller] if isinstance(controller, (str, type)) else list(controller)
        if len(paths) != n_drones:                                  # sim.py:61-62
            paths = (paths * n_drones)[:n_drones]
        classes = [p if isinstance(p, type) else load_controller(p) for p in paths]
    batch_ctrl = all(hasattr(c, "predict_batch") for c in classes) and len(set(classes)) == 1
    env = MultiRaceAviary(cfg, num_drones=n_drones, physics=physics, racemode=racemode, num_envs=max(n_runs, 2),
                          device=device, precision=precision, auto_reset=False, seed=seed or 0,
                          command_actions=not batch_ctrl)
    N, M, ctrl_freq = env.NUM_ENVS, n_drones, cfg["ctrl_freq"]
    import torch
    obs, info = env.reset()
    agents = None
    if not batch_ctrl:
        obs_h = obs.cpu().numpy()
        agents = [[classes[i](i, obs_h[r, i], dict(info, delay=i)) for i in range(M)] for r in range(n_runs)]
    logger = Logger(logging_freq_hz=ctrl_freq, output_folder=output_folder, num_drones=M) if log_env is not None else None
    done = torch.zeros(N, dtype=torch.bool, device=obs.device)
    done[n_runs:] = True                                            # padding env when n_runs == 1
    times = torch.zeros(N, dtype=torch.float64, device=obs.device)
    rewards = torch.zeros(N, dtype=torch.float64, device=obs.device)
    max_steps = int(cfg["episode_len_sec"] * ctrl_freq) + 2
    actions = torch.zeros((N, M, env.handle.A), dtype=obs.dtype, device=obs.device)
    step = 0
    while not bool(done.all()) and step <= max_steps:
        sim_time = step / ctrl_freq                                 # sim.py:79
        if batch_ctrl:
            actions = classes[0].predict_batch(obs, ep_time=sim_time)
        else:
            obs_h, done_h = obs.cpu().numpy(), done.cpu().numpy()
            rows = actions.cpu().numpy()
            for r in range(n_runs):
                if not done_h[r]:
                    for i in range(M):
                        rows[r, i] = _row(agents[r][i].predict(obs_h[r, i], ep_time=sim_time), env.step_counter)
            actions = torch.as_tensor(rows, dtype=obs.dtype, device=obs.device)
        if logger is not None and not bool(done[log_env]):
            state = env.handle.get_state()[log_env].cpu().numpy()
            for i in range(M):
                logger.log(drone=i, timestamp=sim_time, state=state[i])
        obs, reward, terminated, truncated, _ = env.step(actions)
        live = ~done
        rewards += torch.where(live, reward.double(), torch.zeros_like(rewards))
        times = torch.where(live, torch.full_like(times, sim_time), times)   # sim.py:104: time of the last step taken
        done |= terminated | truncated
        step += 1
    if logger is not None:
        logger.save()
    env.close()
    out = times[:n_runs].cpu().tolist()
    if return_stats:
        return {"episode_times": out, "episode_rewards": rewards[:n_runs].cpu().tolist(), "control_steps": step}
    return out


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--config", default="config/getting_started.yaml")
    ap.add_argument("--controller", nargs="*", default=None)
    ap.add_argument("--n_runs", type=int, default=10)
    ap.add_argument("--n_drones", type=int, default=2)
    ap.add_argument("--racemode", default="COMPARE", choices=["COMPARE", "COMPETE"])
    ap.add_argument("--precision", default="fp32", choices=["fp32", "fp64"])
    ap.add_argument("--log_env", type=int, default=None)
    a = ap.parse_args(argv)
    times = simulate(a.config, a.controller, a.n_runs, a.n_drones, racemode=RaceMode[a.racemode], precision=a.precision,
                     log_env=a.log_env)
    print(json.dumps({"episode_times": times if len(times) <= 32 else times[:32] + ["..."],
                      "mean_episode_time": float(np.mean(times))}))


if __name__ == "__main__":
    main()
