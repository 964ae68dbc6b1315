"""examples/downwash.py: two quadrotors on the same x axis at z = 1 and z = 0.5 sweep past each other under
``Physics.PYB_DW``, so the lower one flies through the upper one's downwash twice per period
(gym_pybullet_adrp/examples/downwash.py:30-123).  Same ``run(...)`` keywords and flags, plus ``--num_envs``.

In this fork the DSLPID loops are unstable with the default CF2X (cf2x_IROS.urdf): the drones tumble within
seconds, in the reference as here; ``--drone cf2p`` holds its heights.

    python -m gym_pybullet_adrp_b200.examples.downwash --duration_sec 12
"""
from __future__ import annotations

import argparse

import numpy as np

from ..control.DSLPIDControl import DSLPIDControl
from ..envs.CtrlAviary import CtrlAviary
from ..utils.enums import DroneModel, Physics
from ..utils.utils import str2bool
from ._loop import fly, refuse_graphics

DEFAULT_DRONE = DroneModel.CF2X      # (the reference writes DroneModel("cf2x"), which its own enum -- "cf2x_IROS" in this fork -- rejects)
DEFAULT_SIMULATION_FREQ_HZ = 240
DEFAULT_CONTROL_FREQ_HZ = 48
DEFAULT_DURATION_SEC = 12
DEFAULT_OUTPUT_FOLDER = 'results'


def run(drone=DEFAULT_DRONE, gui=False, record_video=False, simulation_freq_hz=DEFAULT_SIMULATION_FREQ_HZ,
        control_freq_hz=DEFAULT_CONTROL_FREQ_HZ, duration_sec=DEFAULT_DURATION_SEC, output_folder=DEFAULT_OUTPUT_FOLDER,
        plot=False, colab=False, num_envs=1, precision="fp64", verbose=True):
    refuse_graphics(gui, record_video, plot)
    N = int(num_envs)
    init_xyzs = np.array([[.5, 0, 1], [-.5, 0, .5]])
    env = CtrlAviary(drone_model=drone, num_drones=2, initial_xyzs=init_xyzs, physics=Physics.PYB_DW, neighbourhood_radius=10,
                     pyb_freq=simulation_freq_hz, ctrl_freq=control_freq_hz, obstacles=True, num_envs=N, precision=precision,
                     auto_reset=False)
    # x = 0.5 cos(2 pi t / 5 s), y = 0; the second drone half a period behind (:57-63)
    num_wp = control_freq_hz * 5
    target_x = 0.5 * np.cos(2 * np.pi * np.arange(num_wp) / num_wp)
    wp = np.array([0, num_wp // 2])
    ctrl = DSLPIDControl(drone_model=drone, n=2 * N, precision=precision)

    def targets(i):
        return np.stack([target_x[(wp + i) % num_wp], np.zeros(2), init_xyzs[:, 2]], axis=1)        # (2, 3)

    def policy(i, obs):
        rpm, _, _ = ctrl.computeControlFromState(control_timestep=env.CTRL_TIMESTEP, state=obs.reshape(2 * N, 20),
                                                 target_pos=np.tile(targets(i), (N, 1)))
        return np.asarray(rpm).reshape((N, 2, 4) if N > 1 else (2, 4))

    logger, obs = fly(env, policy, lambda i, j: np.hstack([targets(i + 1)[j], np.zeros(9)]), int(duration_sec * env.CTRL_FREQ),
                      output_folder=output_folder, colab=colab, csv_name="dw", gui=gui, verbose=verbose)
    env.close()
    return logger, obs


def main(argv=None):
    parser = argparse.ArgumentParser(description='Downwash example script using CtrlAviary and DSLPIDControl')
    parser.add_argument('--drone', default=DEFAULT_DRONE, type=DroneModel, help='Drone model (default: CF2X)', metavar='', choices=DroneModel)
    parser.add_argument('--gui', default=False, type=str2bool, help='Whether to use PyBullet GUI (unavailable on this box)', metavar='')
    parser.add_argument('--record_video', default=False, type=str2bool, help='Whether to record a video (unavailable on this box)', metavar='')
    parser.add_argument('--simulation_freq_hz', default=DEFAULT_SIMULATION_FREQ_HZ, type=int, help='Simulation frequency in Hz (default: 240)', metavar='')
    parser.add_argument('--control_freq_hz', default=DEFAULT_CONTROL_FREQ_HZ, type=int, help='Control frequency in Hz (default: 48)', metavar='')
    parser.add_argument('--duration_sec', default=DEFAULT_DURATION_SEC, type=int, help='Duration of the simulation in seconds (default: 12)', metavar='')
    parser.add_argument('--output_folder', default=DEFAULT_OUTPUT_FOLDER, type=str, help='Folder where to save logs (default: "results")', metavar='')
    parser.add_argument('--plot', default=False, type=str2bool, help='Whether to plot the simulation results (unavailable on this box)', metavar='')
    parser.add_argument('--colab', default=False, type=bool, help='Whether example is being run by a notebook (default: "False")', metavar='')
    parser.add_argument('--num_envs', default=1, type=int, help='Copies of the scenario flown side by side (default: 1)', metavar='')
    parser.add_argument('--precision', default="fp64", choices=["fp32", "fp64"], help='Build of the library (default: fp64)', metavar='')
    run(**vars(parser.parse_args(argv)))


if __name__ == "__main__":
    main()
