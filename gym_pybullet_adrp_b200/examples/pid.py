"""examples/pid.py: ``num_drones`` quadrotors circle around (0, -0.3) at staggered heights, CtrlAviary + DSLPIDControl
(gym_pybullet_adrp/examples/pid.py:49-177).  Same ``run(...)`` keywords and flags; ``--num_envs N`` flies N copies at once
(one ``DSLPIDControl`` object holds the integrators of all N x num_drones drones, one launch per step for all of them).

In this fork the DSLPID loops are unstable with the default CF2X (cf2x_IROS.urdf): the drones tumble within
seconds, in the reference as here; ``--drone cf2p`` holds its heights.

    python -m gym_pybullet_adrp_b200.examples.pid --num_drones 3 --duration_sec 12 --num_envs 4096
"""
from __future__ import annotations

import argparse

import numpy as np

from ..control.DSLPIDControl import DSLPIDControl
from ..envs.CtrlAviary import CtrlAviary
from ..utils.enums import DroneModel, Physics
from ..utils.utils import str2bool
from ._loop import fly, refuse_graphics

DEFAULT_DRONES = DroneModel.CF2X     # (the reference writes DroneModel("cf2x"), which its own enum -- "cf2x_IROS" in this fork -- rejects)
DEFAULT_NUM_DRONES = 3
DEFAULT_PHYSICS = Physics("pyb")
DEFAULT_SIMULATION_FREQ_HZ = 240
DEFAULT_CONTROL_FREQ_HZ = 48
DEFAULT_DURATION_SEC = 12
DEFAULT_OUTPUT_FOLDER = 'results'


def run(drone=DEFAULT_DRONES, num_drones=DEFAULT_NUM_DRONES, physics=DEFAULT_PHYSICS, gui=False, record_video=False, plot=False,
        user_debug_gui=False, obstacles=True, simulation_freq_hz=DEFAULT_SIMULATION_FREQ_HZ,
        control_freq_hz=DEFAULT_CONTROL_FREQ_HZ, duration_sec=DEFAULT_DURATION_SEC, output_folder=DEFAULT_OUTPUT_FOLDER,
        colab=False, num_envs=1, precision="fp64", verbose=True):
    refuse_graphics(gui, record_video, plot)
    M, N = int(num_drones), int(num_envs)
    height, height_step, radius = .1, .05, .3
    phase = (np.arange(M) / 6) * 2 * np.pi + np.pi / 2
    init_xyzs = np.stack([radius * np.cos(phase), radius * np.sin(phase) - radius, height + np.arange(M) * height_step], axis=1)
    init_rpys = np.stack([np.zeros(M), np.zeros(M), np.arange(M) * (np.pi / 2) / M], axis=1)
    # one lap of the circle in 10 s (:71-77); drone i starts a sixth of a lap after drone i - 1
    num_wp = control_freq_hz * 10
    lap = (np.arange(num_wp) / num_wp) * 2 * np.pi + np.pi / 2
    target_xy = np.stack([radius * np.cos(lap) + init_xyzs[0, 0], radius * np.sin(lap) - radius + init_xyzs[0, 1]], axis=1)
    wp = np.array([int((i * num_wp / 6) % num_wp) for i in range(M)])

    env = CtrlAviary(drone_model=drone, num_drones=M, initial_xyzs=init_xyzs, initial_rpys=init_rpys, physics=physics,
                     neighbourhood_radius=10, pyb_freq=simulation_freq_hz, ctrl_freq=control_freq_hz, obstacles=obstacles,
                     user_debug_gui=user_debug_gui, num_envs=N, precision=precision, auto_reset=False)
    ctrl = DSLPIDControl(drone_model=drone, n=N * M, precision=precision)

    def targets(i):
        k = (wp + i) % num_wp
        return np.concatenate([target_xy[k], init_xyzs[:, 2:3]], axis=1)                     # (M, 3)

    def policy(i, obs):
        rows = obs.reshape(N * M, 20)
        rpm, _, _ = ctrl.computeControlFromState(control_timestep=env.CTRL_TIMESTEP, state=rows,
                                                 target_pos=np.tile(targets(i), (N, 1)), target_rpy=np.tile(init_rpys, (N, 1)))
        return np.asarray(rpm).reshape((N, M, 4) if N > 1 else (M, 4))

    def control_row(i, j):
        return np.hstack([targets(i + 1)[j], init_rpys[j], np.zeros(6)])                     # (logged after the counter moved, :150-156)

    logger, obs = fly(env, policy, control_row, int(duration_sec * env.CTRL_FREQ), output_folder=output_folder, colab=colab,
                      csv_name="pid", gui=gui, verbose=verbose)
    env.close()
    return logger, obs


def main(argv=None):
    parser = argparse.ArgumentParser(description='Helix flight script using CtrlAviary and DSLPIDControl')
    parser.add_argument('--drone', default=DEFAULT_DRONES, type=DroneModel, help='Drone model (default: CF2X)', metavar='', choices=DroneModel)
    parser.add_argument('--num_drones', default=DEFAULT_NUM_DRONES, type=int, help='Number of drones (default: 3)', metavar='')
    parser.add_argument('--physics', default=DEFAULT_PHYSICS, type=Physics, help='Physics updates (default: PYB)', metavar='', choices=Physics)
    parser.add_argument('--gui', default=False, type=str2bool, help='Whether to use PyBullet GUI (unavailable on this box)', metavar='')
    parser.add_argument('--record_video', default=False, type=str2bool, help='Whether to record a video (unavailable on this box)', metavar='')
    parser.add_argument('--plot', default=False, type=str2bool, help='Whether to plot the simulation results (unavailable on this box)', metavar='')
    parser.add_argument('--user_debug_gui', default=False, type=str2bool, help='Whether to add debug lines and parameters to the GUI', metavar='')
    parser.add_argument('--obstacles', default=True, type=str2bool, help='Whether to add obstacles to the environment (default: True)', metavar='')
    parser.add_argument('--simulation_freq_hz', default=DEFAULT_SIMULATION_FREQ_HZ, type=int, help='Simulation frequency in Hz (default: 240)', metavar='')
    parser.add_argument('--control_freq_hz', default=DEFAULT_CONTROL_FREQ_HZ, type=int, help='Control frequency in Hz (default: 48)', metavar='')
    parser.add_argument('--duration_sec', default=DEFAULT_DURATION_SEC, type=int, help='Duration of the simulation in seconds (default: 12)', metavar='')
    parser.add_argument('--output_folder', default=DEFAULT_OUTPUT_FOLDER, type=str, help='Folder where to save logs (default: "results")', metavar='')
    parser.add_argument('--colab', default=False, type=bool, help='Whether example is being run by a notebook (default: "False")', metavar='')
    parser.add_argument('--num_envs', default=1, type=int, help='Copies of the scenario flown side by side (default: 1)', metavar='')
    parser.add_argument('--precision', default="fp64", choices=["fp32", "fp64"], help='Build of the library (default: fp64, the reference\'s numpy)', metavar='')
    run(**vars(parser.parse_args(argv)))


if __name__ == "__main__":
    main()
