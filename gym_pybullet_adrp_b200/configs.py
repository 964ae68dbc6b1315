"""Race-track configurations (the reference's config/level*.yaml, SURVEY section 5 'Config / flags').

The six reference files differ only in which randomisation blocks are switched on and, for
``twogates``, in the track itself (config/level3.yaml:13-111, config/twogates.yaml:13-57); they are
generated here from one table so that the GPU box needs no file from the reference.  ``load_config``
also reads any YAML file with the same keys (the reference's ``utils.load_config``, utils/utils.py:95-108).
"""
from __future__ import annotations

import copy
from pathlib import Path


class Config(dict):
    """Attribute-access dict (stand-in for ``munch.Munch``; same access pattern as the reference)."""
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__

    def __deepcopy__(self, memo):
        return Config((k, copy.deepcopy(v, memo)) for k, v in self.items())


def configify(x):
    if isinstance(x, dict):
        return Config((k, configify(v)) for k, v in x.items())
    if isinstance(x, list):
        return [configify(v) for v in x]
    return x


_U = "uniform"
_BLOCKS = {
    "random_drone_state_info": {"pos": {"distrib": _U, "x": [-0.1, 0.1], "y": [-0.1, 0.1], "z": [0.0, 0.02]},
                                "rot": {"distrib": _U, "r": [-0.1, 0.1], "p": [-0.1, 0.1], "y": [-0.1, 0.1]}},
    "random_drone_inertia_info": {"M": {"distrib": _U, "range": [-0.01, 0.01]},
                                  "Ixx": {"distrib": _U, "range": [-1e-6, 1e-6]},
                                  "Iyy": {"distrib": _U, "range": [-1e-6, 1e-6]},
                                  "Izz": {"distrib": _U, "range": [-1e-6, 1e-6]}},
    "random_gates_obstacles_info": {"gates": {"distrib": _U, "range": [-0.15, 0.15]},
                                    "obstacles": {"distrib": _U, "range": [-0.15, 0.15]}},
    "disturbances_info": {"action": {"distrib": "normal", "std": 0.001},
                          "dynamics": {"distrib": _U, "low": [-0.1, -0.1, -0.1], "high": [0.1, 0.1, 0.1]}},
}
_SWITCHES = ["random_drone_state", "random_drone_inertia", "random_gates_obstacles", "disturbances"]
_LEVEL_FLAGS = {"getting_started": (0, 0, 0, 0), "level0": (1, 0, 0, 0), "level1": (1, 1, 0, 1),
                "level2": (1, 1, 1, 1), "level3": (1, 1, 1, 1), "twogates": (1, 0, 0, 0)}
_TRACK = {
    "gates": [[0.45, -1.0, 0.525, 0, 0, 2.35, 1], [1.0, -1.55, 1.0, 0, 0, -0.78, 0],
              [0.0, 0.5, 0.525, 0, 0, 0, 1], [-0.5, -0.5, 1.0, 0, 0, 3.14, 0]],
    "obstacles": [[1.0, -0.5, 0.525, 0, 0, 0], [0.5, -1.5, 0.525, 0, 0, 0], [-0.5, 0, 0.525, 0, 0, 0],
                  [0, 1.0, 0.525, 0, 0, 0]],
    "starts": [[0.9, 0.9, 0.05], [1.1, 1.1, 0.05]],
}
_TRACK_TWOGATES = {
    "gates": [[1.0, 0.4, 0.525, 0, 0, 0, 1.0], [1.0, -1.0, 1.0, 0, 0, 0, 0], [5.0, -2.7, 1.0, 0, 0, 0, 0],
              [5.0, -2.0, 1.0, 0, 0, 0, 0]],
    "obstacles": [[3.0, -0.5, 0.525, 0, 0, 0], [3, -1.5, 0.525, 0, 0, 0], [7, 0, 0.525, 0, 0, 0], [5, 0, 0.525, 0, 0, 0]],
    "starts": [[0.9, 1.0, 0.05], [1.1, 1.0, 0.05]],
}


def builtin_config(name: str) -> Config:
    if name not in _LEVEL_FLAGS:
        raise KeyError(f"unknown built-in race config {name!r}; have {sorted(_LEVEL_FLAGS)}")
    trk = _TRACK_TWOGATES if name == "twogates" else _TRACK
    cfg = {
        "seed": 1337, "reseed_on_reset": name != "level3", "camera_view": [5, -40, -40, 0.5, -1, 0.5],
        "bounds": [[-3, -3, 0], [3, 3, 2]],
        "ctrl_freq": 30 if name == "getting_started" else 25, "pyb_freq": 500, "episode_len_sec": 33,
        "done_on_completion": True, "done_on_collision": True,
        "init_states": {f"drone{i}": {"pos": list(p), "vel": [0, 0, 0], "rpy": [0, 0, 0], "pqr": [0, 0, 0]}
                        for i, p in enumerate(trk["starts"])},
        "gates": copy.deepcopy(trk["gates"]), "obstacles": copy.deepcopy(trk["obstacles"]),
    }
    for key, on in zip(_SWITCHES, _LEVEL_FLAGS[name]):
        cfg[key] = bool(on)
        if on:
            cfg[key + "_info"] = copy.deepcopy(_BLOCKS[key + "_info"])
    return configify(cfg)


def load_config(path) -> Config:
    """``load_config("level3")`` (built-in) or ``load_config("path/to/level3.yaml")`` (utils/utils.py:95-108)."""
    if isinstance(path, str) and path in _LEVEL_FLAGS:
        return builtin_config(path)
    import yaml
    path = Path(path)
    assert path.exists(), f"Configuration file not found: {path}"
    with open(path, "r", encoding="utf-8") as file:
        return configify(yaml.safe_load(file))
