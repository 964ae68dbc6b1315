"""examples/pid_velocity.py: four quadrotors follow piecewise-constant velocity commands through ``VelocityAviary``
(gym_pybullet_adrp/examples/pid_velocity.py:48-155): out along +y, then back.  Same ``run(...)`` keywords and flags, plus
``--num_envs``.

In this fork the DSLPID loops are unstable with the default CF2X (cf2x_IROS.urdf): the drones tumble within
seconds, in the reference as here; ``--drone cf2p`` holds its heights.

    python -m gym_pybullet_adrp_b200.examples.pid_velocity --duration_sec 5
"""
from __future__ import annotations

import argparse

import numpy as np

from ..envs.VelocityAviary import VelocityAviary
from ..utils.enums import DroneModel, Physics
from ..utils.utils import str2bool
from ._loop import fly, refuse_graphics

DEFAULT_DRONE = DroneModel.CF2X      # (the reference writes DroneModel("cf2x"), which its own enum -- "cf2x_IROS" in this fork -- rejects)
DEFAULT_SIMULATION_FREQ_HZ = 240
DEFAULT_CONTROL_FREQ_HZ = 48
DEFAULT_DURATION_SEC = 5
DEFAULT_OUTPUT_FOLDER = 'results'


def run(drone=DEFAULT_DRONE, gui=False, record_video=False, plot=False, user_debug_gui=False, obstacles=False,
        simulation_freq_hz=DEFAULT_SIMULATION_FREQ_HZ, control_freq_hz=DEFAULT_CONTROL_FREQ_HZ, duration_sec=DEFAULT_DURATION_SEC,
        output_folder=DEFAULT_OUTPUT_FOLDER, colab=False, num_envs=1, precision="fp64", verbose=True):
    refuse_graphics(gui, record_video, plot)
    N = int(num_envs)
    init_xyzs = np.array([[0, 0, .1], [.3, 0, .1], [.6, 0, .1], [0.9, 0, .1]])
    init_rpys = np.array([[0, 0, 0], [0, 0, np.pi / 3], [0, 0, np.pi / 4], [0, 0, np.pi / 2]])
    env = VelocityAviary(drone_model=drone, num_drones=4, initial_xyzs=init_xyzs, initial_rpys=init_rpys, physics=Physics.PYB,
                         neighbourhood_radius=10, pyb_freq=simulation_freq_hz, ctrl_freq=control_freq_hz, obstacles=obstacles,
                         user_debug_gui=user_debug_gui, num_envs=N, precision=precision, auto_reset=False)
    # [direction xyz, fraction of the speed limit]; each drone turns back a sixth of the flight after the previous one (:97-102)
    num_wp = control_freq_hz * duration_sec
    out = np.array([[-0.5, 1, 0, 0.99], [0, 1, 0, 0.99], [0.2, 1, 0.2, 0.99], [0, 1, 0.5, 0.99]])
    back = out * np.array([-1, -1, -1, 1])
    turn = num_wp / 8 + np.arange(4) * num_wp / 6
    target_vel = np.where((np.arange(num_wp)[None, :] < turn[:, None])[..., None], out[:, None, :], back[:, None, :])   # (4, num_wp, 4)

    def command(i):
        return target_vel[np.arange(4), i % num_wp]                                          # (4, 4)

    def policy(i, obs):
        return np.broadcast_to(command(i), (N, 4, 4)).copy() if N > 1 else command(i).copy()

    logger, obs = fly(env, policy, lambda i, j: np.hstack([command(i + 1)[j, 0:3], np.zeros(9)]), int(duration_sec * env.CTRL_FREQ),
                      output_folder=output_folder, colab=colab, csv_name="vel", gui=gui, verbose=verbose)
    env.close()
    return logger, obs


def main(argv=None):
    parser = argparse.ArgumentParser(description='Velocity control example using VelocityAviary')
    parser.add_argument('--drone', default=DEFAULT_DRONE, type=DroneModel, help='Drone model (default: CF2X)', metavar='', choices=DroneModel)
    parser.add_argument('--gui', default=False, type=str2bool, help='Whether to use PyBullet GUI (unavailable on this box)', metavar='')
    parser.add_argument('--record_video', default=False, type=str2bool, help='Whether to record a video (unavailable on this box)', metavar='')
    parser.add_argument('--plot', default=False, type=str2bool, help='Whether to plot the simulation results (unavailable on this box)', metavar='')
    parser.add_argument('--user_debug_gui', default=False, type=str2bool, help='Whether to add debug lines and parameters to the GUI', metavar='')
    parser.add_argument('--obstacles', default=False, type=str2bool, help='Whether to add obstacles to the environment (default: False)', metavar='')
    parser.add_argument('--simulation_freq_hz', default=DEFAULT_SIMULATION_FREQ_HZ, type=int, help='Simulation frequency in Hz (default: 240)', metavar='')
    parser.add_argument('--control_freq_hz', default=DEFAULT_CONTROL_FREQ_HZ, type=int, help='Control frequency in Hz (default: 48)', metavar='')
    parser.add_argument('--duration_sec', default=DEFAULT_DURATION_SEC, type=int, help='Duration of the simulation in seconds (default: 5)', metavar='')
    parser.add_argument('--output_folder', default=DEFAULT_OUTPUT_FOLDER, type=str, help='Folder where to save logs (default: "results")', metavar='')
    parser.add_argument('--colab', default=False, type=bool, help='Whether example is being run by a notebook (default: "False")', metavar='')
    parser.add_argument('--num_envs', default=1, type=int, help='Copies of the scenario flown side by side (default: 1)', metavar='')
    parser.add_argument('--precision', default="fp64", choices=["fp32", "fp64"], help='Build of the library (default: fp64)', metavar='')
    run(**vars(parser.parse_args(argv)))


if __name__ == "__main__":
    main()
