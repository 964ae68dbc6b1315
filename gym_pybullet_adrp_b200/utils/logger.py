"""``Logger`` (gym_pybullet_adrp/utils/logger.py:9-129): same arrays, same ``.npy`` (npz) layout, so flights dumped
from a batched run open in the reference's own plotting tools.  ``plot()`` needs matplotlib and a display:
unavailable on this box."""
from __future__ import annotations

import os
from datetime import datetime

import numpy as np


class Logger:
    def __init__(self, logging_freq_hz: int, output_folder: str = "results", num_drones: int = 1, duration_sec: int = 0,
                 colab: bool = False):
        self.COLAB, self.OUTPUT_FOLDER = colab, output_folder
        os.makedirs(self.OUTPUT_FOLDER, exist_ok=True)
        self.LOGGING_FREQ_HZ, self.NUM_DRONES = logging_freq_hz, num_drones
        self.PREALLOCATED_ARRAYS = duration_sec != 0
        T = duration_sec * self.LOGGING_FREQ_HZ
        self.counters = np.zeros(num_drones)
        self.timestamps = np.zeros((num_drones, T))
        self.states = np.zeros((num_drones, 16, T))       # pos3 vel3 rpy3 ang_vel3 rpm4 (logger.py:43-58)
        self.controls = np.zeros((num_drones, 12, T))     # target pos3 vel3 rpy3 ang_vel3 (:60-71)

    def log(self, drone: int, timestamp, state, control=np.zeros(12)):
        """logger.py:75-113.  ``state``: the 20-float row of ``_getDroneStateVector``."""
        if drone < 0 or drone >= self.NUM_DRONES or timestamp < 0 or len(state) != 20 or len(control) != 12:
            print("[ERROR] in Logger.log(), invalid data")
        k = int(self.counters[drone])
        if k >= self.timestamps.shape[1]:
            self.timestamps = np.concatenate((self.timestamps, np.zeros((self.NUM_DRONES, 1))), axis=1)
            self.states = np.concatenate((self.states, np.zeros((self.NUM_DRONES, 16, 1))), axis=2)
            self.controls = np.concatenate((self.controls, np.zeros((self.NUM_DRONES, 12, 1))), axis=2)
        elif not self.PREALLOCATED_ARRAYS and self.timestamps.shape[1] > k:
            k = self.timestamps.shape[1] - 1
        self.timestamps[drone, k] = timestamp
        state = np.asarray(state)
        self.states[drone, :, k] = np.hstack([state[0:3], state[10:13], state[7:10], state[13:20]])
        self.controls[drone, :, k] = control
        self.counters[drone] = k + 1

    def save(self):
        """logger.py:117-121: one ``np.savez`` archive named ``save-flight-<date>.npy``."""
        path = os.path.join(self.OUTPUT_FOLDER, "save-flight-" + datetime.now().strftime("%m.%d.%Y_%H.%M.%S") + ".npy")
        with open(path, "wb") as out_file:
            np.savez(out_file, timestamps=self.timestamps, states=self.states, controls=self.controls)
        return path

    def save_as_csv(self, comment: str = ""):
        """logger.py:131-201: a folder ``save-flight-<comment>-<date>`` with one two-column (time, value) CSV per logged
        quantity and drone -- x y z, r p ya (+ their finite-difference rates rr pr yar), vx vy vz, wx wy wz, rpm0-3, pwm0-3."""
        csv_dir = os.path.join(self.OUTPUT_FOLDER, "save-flight-" + comment + "-" + datetime.now().strftime("%m.%d.%Y_%H.%M.%S"))
        os.makedirs(csv_dir, exist_ok=True)
        t = np.arange(0, self.timestamps.shape[1] / self.LOGGING_FREQ_HZ, 1 / self.LOGGING_FREQ_HZ)[:self.timestamps.shape[1]]
        rows = {"x": 0, "y": 1, "z": 2, "r": 6, "p": 7, "ya": 8, "vx": 3, "vy": 4, "vz": 5, "wx": 9, "wy": 10, "wz": 11}
        for i in range(self.NUM_DRONES):
            cols = {name + str(i): self.states[i, k, :] for name, k in rows.items()}
            for name, k in (("rr", 6), ("pr", 7), ("yar", 8)):             # rates of the logged angles (:161-169)
                cols[name + str(i)] = np.hstack([0, np.diff(self.states[i, k, :]) * self.LOGGING_FREQ_HZ])
            for m in range(4):
                cols[f"rpm{m}-{i}"] = self.states[i, 12 + m, :]
                cols[f"pwm{m}-{i}"] = (self.states[i, 12 + m, :] - 4070.3) / 0.2685
            for name, v in cols.items():
                with open(os.path.join(csv_dir, name + ".csv"), "wb") as out_file:
                    np.savetxt(out_file, np.transpose(np.vstack([t, v])), delimiter=",")
        return csv_dir

    def plot(self, pwm=False):
        raise NotImplementedError("Logger.plot needs matplotlib and a display: unavailable on this box")
