"""Golden vectors for the race env: runs the REFERENCE'S OWN ``MultiRaceAviary`` + ``MellingerControl``
(one controller process per drone over mp.Pipe, exactly as envs/MultiRaceAviary.py:107-115 does) with
the stand-ins of oracle/refstub for PyBullet and for the pycffirmware SWIG module.

    python tests/golden/make_golden_race.py        (build container only; needs /root/reference)

Pinned by these vectors: every reference-owned line of MultiRaceAviary.step / reset / _gate_progress /
_computeObs / _computeTerminated and of the MellingerControl wrapper (sensor synthesis, sub-rate
schedule, power distribution, battery curve, thrust->PWM->RPM, prop reorder).  NOT pinned (declared
restatements): the firmware core, Bullet's integrator and collision / ray queries.
The reference's RNG is unseeded; every draw it makes is recorded here and stored next to
the outputs so that the oracle and the CUDA path can be fed the same numbers.
"""
from __future__ import annotations

import json
import os
import sys
import warnings

warnings.filterwarnings("ignore")
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.environ.get("ADRP_GOLDEN_OUT", HERE)      # tests/test_golden_regen.py regenerates into a scratch directory
REF = os.environ.get("ADRP_REFERENCE_ROOT", "/root/reference")
sys.path.insert(0, os.path.join(ROOT, "oracle", "refstub"))
sys.path.insert(0, REF)
sys.path.insert(0, ROOT)
os.chdir(REF)                      # the reference opens "gym_pybullet_adrp/assets/..." relative to its root

import numpy as np  # noqa: E402

import firmware_stub  # noqa: E402
import gym_pybullet_adrp.control  # noqa: E402,F401
from gym_pybullet_adrp.envs.MultiRaceAviary import MultiRaceAviary  # noqa: E402
from gym_pybullet_adrp.utils.enums import DroneModel, Physics, RaceMode  # noqa: E402
from munch import munchify  # noqa: E402

from gym_pybullet_adrp_b200 import configs  # noqa: E402

# pycffirmware is not installable here; the package re-exports the class under the submodule's name, hence sys.modules
sys.modules["gym_pybullet_adrp.control.MellingerControl"].load_firmware = lambda path: firmware_stub.Firmware()


class RecordingRNG:
    """Stands in for np.random / env.np_random: same formulas (low + (high-low)*u, N(0, std)), every unit
    draw logged in call order."""

    def __init__(self, seed):
        self.g = np.random.default_rng(seed)
        self.uniforms, self.step_noise = [], []

    def uniform(self, low=0.0, high=1.0, size=None):
        low, high = np.asarray(low, float), np.asarray(high, float)
        shape = size if size is not None else np.broadcast(low, high).shape
        u = self.g.random(shape)
        self.uniforms.extend(np.ravel(u).tolist())
        return low + (high - low) * u

    def normal(self, loc=0.0, scale=1.0, size=None):
        v = loc + scale * self.g.standard_normal(size)
        self.step_noise.append(("n", np.array(v)))
        return v


def gate_pass_setpoints(cfg, steps, M, through_gate1=False):
    """Smooth full-state setpoints: climb, fly through gate 0 along its normal, carry on to gate 1
    (through_gate1: around the obstacle next to it and through gate 1 as well)."""
    g0, g1 = cfg.gates[0], cfg.gates[1]
    n0 = np.array([-np.sin(g0[5]), np.cos(g0[5])])
    n1 = np.array([-np.sin(g1[5]), np.cos(g1[5])])
    starts = [np.array(cfg.init_states[f"drone{i}"].pos) for i in range(M)]
    out = np.zeros((steps, M, 4))
    for i in range(M):
        way = [np.r_[starts[i][:2], 0.35], np.r_[1.0, 0.0, 0.45], np.r_[np.array(g0[:2]) - 0.45 * n0, 0.525],
               np.r_[np.array(g0[:2]) + 0.45 * n0, 0.525], np.r_[g1[0] - 0.3, g1[1] - 0.1, 0.9]]
        seg_t = [1.2, 2.2, 2.0, 1.4, 1.8]
        if through_gate1:
            way = way[:4] + [np.r_[0.15, -1.75, 0.9], np.r_[np.array(g1[:2]) - 0.45 * n1, 1.0],
                             np.r_[np.array(g1[:2]) + 0.45 * n1, 1.0]]
            seg_t = [1.2, 2.2, 2.0, 1.4, 1.6, 1.8, 1.6]
        t0 = 1.6 * i                                   # the second drone follows later
        for k in range(steps):
            t = max(0.0, k / 25.0 - t0)
            p, acc = np.r_[starts[i][:2], 0.05], 0.0
            for w, dt in zip(way, seg_t):
                if t <= acc + dt:
                    s = (t - acc) / dt
                    s = s * s * (3 - 2 * s)
                    p = p + s * (w - p)
                    break
                acc += dt
                p = w
            out[k, i, :3] = p
            out[k, i, 3] = 0.0
    return out


def race_case(name, level, physics, racemode, M, steps, seed, actions=None):
    cfg = munchify(json.loads(json.dumps(configs.builtin_config(level))))
    rec = RecordingRNG(seed)
    orig_uniform = np.random.uniform
    np.random.uniform = rec.uniform                              # _addObstacles uses the global np.random (:362,373)
    env = MultiRaceAviary(cfg, drone_model=DroneModel.CF2X, num_drones=M, physics=physics, racemode=racemode)
    env.np_random = rec
    try:
        obs0, _ = env.reset()
        draws = np.array(rec.uniforms)
        state0 = np.array([env._getDroneStateVector(i) for i in range(M)])
        for i in range(M):                                       # pose AFTER _drone_init (initial_obs predates it, E17)
            env._updateAndStoreKinematicInformation()
        state0 = np.array([env._getDroneStateVector(i) for i in range(M)])
        if actions is None:
            actions = gate_pass_setpoints(cfg, steps, M)
        S = env.PYB_STEPS_PER_CTRL
        noise = np.zeros((steps, M, S, 7))
        obs = np.zeros((steps,) + obs0.shape)
        term, trunc = np.zeros(steps, bool), np.zeros(steps, bool)
        gate, elim, fin = np.zeros((steps, M)), np.zeros((steps, M), bool), np.zeros((steps, M), bool)
        for k in range(steps):
            rec.uniforms, rec.step_noise = [], []
            o, r, te, tr, _ = env.step(actions[k].copy())
            obs[k], term[k], trunc[k] = o, te, tr
            gate[k], elim[k], fin[k] = env.current_gate, env.drones_eliminated, env.drones_finished
            if cfg.disturbances:
                u = np.array(rec.uniforms).reshape(S, M, 3)
                lo, hi = np.array(cfg.disturbances_info.dynamics.low), np.array(cfg.disturbances_info.dynamics.high)
                noise[k, :, :, 0:3] = np.transpose(lo + (hi - lo) * u, (1, 0, 2))
                nrm = np.array([v for _, v in rec.step_noise])   # (S, M, 4)
                noise[k, :, :, 3:7] = np.transpose(nrm, (1, 0, 2))
    finally:
        np.random.uniform = orig_uniform
        env.close()
    meta = dict(env="race", level=level, physics=physics.name, racemode=racemode.name, M=M, steps=steps)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), meta=json.dumps(meta), draws=draws, state0=state0,
                        obs0=obs0, actions=actions, noise=noise, obs=obs, terminated=term, truncated=trunc,
                        current_gate=gate, eliminated=elim, finished=fin)
    print(name, "gates passed:", gate[-1].tolist(), "eliminated:", elim[-1].tolist(), "terminated at:",
          int(np.argmax(term)) if term.any() else None)


def wrapper_case(name, level, M, steps, seed):
    """DroneObservationWrapper(RewardWrapper(MultiRaceAviary)) as the reference composes them for training
    (utils/wrapper.py): shaped reward of drone 0, yaw zeroed, terminated after two gates."""
    from gym_pybullet_adrp.utils.wrapper import DroneObservationWrapper, RewardWrapper
    cfg = munchify(json.loads(json.dumps(configs.builtin_config(level))))
    rec = RecordingRNG(seed)
    orig_uniform = np.random.uniform
    np.random.uniform = rec.uniform
    base = MultiRaceAviary(cfg, drone_model=DroneModel.CF2X, num_drones=M, physics=Physics.PYB, racemode=RaceMode.COMPARE)
    base.np_random = rec
    env = DroneObservationWrapper(RewardWrapper(base))
    try:
        obs0, _ = env.reset()
        draws = np.array(rec.uniforms)
        actions = gate_pass_setpoints(cfg, steps, M, through_gate1=True)
        actions[:, :, 3] = 0.3                                   # the wrapper must zero this
        obs = np.zeros((steps,) + obs0.shape)
        rew, term, trunc, gate = np.zeros(steps), np.zeros(steps, bool), np.zeros(steps, bool), np.zeros((steps, M))
        n = 0
        for k in range(steps):
            o, r, te, tr, _ = env.step(actions[k].copy())
            obs[k], rew[k], term[k], trunc[k], gate[k] = o, r, te, tr, base.current_gate
            n = k + 1
            if te or tr:
                break
    finally:
        np.random.uniform = orig_uniform
        base.close()
    meta = dict(env="race", level=level, physics="PYB", racemode="COMPARE", M=M, steps=n, wrappers=3)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), meta=json.dumps(meta), draws=draws, obs0=obs0,
                        actions=actions[:n], obs=obs[:n], reward=rew[:n], terminated=term[:n], truncated=trunc[:n],
                        current_gate=gate[:n])
    print(name, "steps:", n, "gates:", gate[n - 1].tolist(), "return:", rew[:n].sum(), "terminated:", term[n - 1])


def fullstate_case(name, level, M, steps, seed):
    """(Command.FULLSTATE, (pos, vel, acc, yaw, rpy_rate, t)) tuples as HardCodedController.predict sends them
    (user_controller/HardCodedController.py:166-176): MultiRaceAviary.step (:198-205) -> MellingerControl.sendFullStateCmd
    (:510-543) with non-zero velocity / acceleration / yaw / body-rate setpoints."""
    from gym_pybullet_adrp.utils.enums import Command
    cfg = munchify(json.loads(json.dumps(configs.builtin_config(level))))
    rec = RecordingRNG(seed)
    orig_uniform = np.random.uniform
    np.random.uniform = rec.uniform
    env = MultiRaceAviary(cfg, drone_model=DroneModel.CF2X, num_drones=M, physics=Physics.PYB, racemode=RaceMode.COMPARE)
    env.np_random = rec
    try:
        obs0, _ = env.reset()
        draws = np.array(rec.uniforms)
        pos = gate_pass_setpoints(cfg, steps + 1, M)[:, :, :3]
        act = np.zeros((steps, M, 13))
        act[:, :, 0:3] = pos[:-1]
        act[:, :, 3:6] = (pos[1:] - pos[:-1]) * 25.0                    # feed-forward velocity of the setpoint path
        act[:, :, 6:9] = 0.5                                            # HardCodedController sends ones(3) * 0.5
        act[:, :, 9] = 0.4 * np.sin(np.arange(steps) / 20.0)[:, None]   # yaw
        act[:, :, 10:13] = 0.05 * np.cos(np.arange(steps) / 15.0)[:, None, None] * np.array([1.0, -1.0, 0.5])
        obs = np.zeros((steps,) + obs0.shape)
        term, trunc, gate = np.zeros(steps, bool), np.zeros(steps, bool), np.zeros((steps, M))
        for k in range(steps):
            cmds = [(Command.FULLSTATE, [act[k, i, 0:3].copy(), act[k, i, 3:6].copy(), act[k, i, 6:9].copy(), float(act[k, i, 9]),
                                         act[k, i, 10:13].copy(), k / 25.0]) for i in range(M)]
            o, r, te, tr, _ = env.step(cmds)
            obs[k], term[k], trunc[k], gate[k] = o, te, tr, env.current_gate
    finally:
        np.random.uniform = orig_uniform
        env.close()
    meta = dict(env="race", level=level, physics="PYB", racemode="COMPARE", M=M, steps=steps, full_state_actions=True)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), meta=json.dumps(meta), draws=draws, obs0=obs0, actions=act,
                        obs=obs, terminated=term, truncated=trunc, current_gate=gate)
    print(name, "gates:", gate[-1].tolist(), "terminated:", term[-1])


CMD_ID = {"non": 0, "fst": 1, "tko": 2, "toy": 3, "tov": 4, "lnd": 5, "ldy": 6, "ldv": 7, "stp": 8, "gto": 9, "ntf": 10}


def command_row(cmd, args):
    """(Command, args) -> the 16-float row of ADRP_ACT_MEL_CMD (include/adrp.h): [command id, flattened args, 0...]."""
    row = np.zeros(16)
    row[0] = CMD_ID[cmd.value]
    flat = np.concatenate([np.atleast_1d(np.asarray(a, dtype=float)) for a in args]) if len(args) else np.zeros(0)
    row[1:1 + len(flat)] = flat
    return row


def _commander_run(name, level, M, steps, seed, policy):
    """The driver loop of scripts/sim.py (:68-100) around the reference's MultiRaceAviary: policy(i, obs_i, sim_time) returns
    what BaseController.predict returns.  Every (Command, args) sent is recorded as a 16-float row."""
    cfg = munchify(json.loads(json.dumps(configs.builtin_config(level))))
    rec = RecordingRNG(seed)
    orig_uniform = np.random.uniform
    np.random.uniform = rec.uniform
    env = MultiRaceAviary(cfg, drone_model=DroneModel.CF2X, num_drones=M, physics=Physics.PYB, racemode=RaceMode.COMPARE)
    env.np_random = rec
    try:
        obs, info = env.reset()
        obs0, draws = obs.copy(), np.array(rec.uniforms)
        agents = policy(obs, info)
        rows = np.zeros((steps, M, 16))
        out = np.zeros((steps,) + obs.shape)
        term, trunc, gate = np.zeros(steps, bool), np.zeros(steps, bool), np.zeros((steps, M))
        n = 0
        for k in range(steps):
            sim_time = k / cfg.ctrl_freq                                  # sim.py:75 (the CONFIG's ctrl_freq)
            actions = [a(obs[i], sim_time) for i, a in enumerate(agents)]
            for i, (cmd, args) in enumerate(actions):
                rows[k, i] = command_row(cmd, args)
            obs, r, te, tr, _ = env.step(actions)
            out[k], term[k], trunc[k], gate[k] = obs, te, tr, env.current_gate
            n = k + 1
            if te or tr:
                break
    finally:
        np.random.uniform = orig_uniform
        env.close()
    meta = dict(env="race", level=level, physics="PYB", racemode="COMPARE", M=M, steps=n, command_actions=True)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), meta=json.dumps(meta), draws=draws, obs0=obs0, actions=rows[:n],
                        obs=out[:n], terminated=term[:n], truncated=trunc[:n], current_gate=gate[:n])
    print(name, "steps:", n, "gates:", gate[n - 1].tolist(), "terminated:", bool(term[n - 1]),
          "commands:", sorted(set(rows[:n, :, 0].astype(int).ravel().tolist())))


def commander_hardcoded_case(name, level, M, steps, seed):
    """user_controller/HardCodedController.py, UNMODIFIED: TAKEOFF [0.3, 2] -> FULLSTATE along its spline -> NOTIFY -> LAND ->
    NONE (:158-188), one instance per drone with info["delay"] = drone id (sim.py:70-72)."""
    from user_controller.HardCodedController import HardCodedController

    def policy(obs, info):
        agents = []
        for i in range(M):
            info = dict(info or {})
            info["delay"] = i
            c = HardCodedController(i, obs[i], info)
            agents.append(lambda o, t, c=c: c.predict(o, ep_time=t))
        return agents
    _commander_run(name, level, M, steps, seed, policy)


def commander_script_case(name, level, M, steps, seed):
    """Every Command of utils/enums.py:58-70 through the reference's low_level_control dispatch (MellingerControl.py:32-57):
    take-off variants, go-to (absolute / relative), the land variants (which the firmware's planner refuses after the
    Stop() that process_command_queue issues first), STOP, NOTIFY, NONE, and FULLSTATE in between."""
    from gym_pybullet_adrp.utils.enums import Command

    def script(i):
        x0 = [1.0, 1.0][i % 2] + 0.0
        def act(o, t, i=i):
            k = int(round(t * 30))
            p = np.array(o[0:3])
            if i % 2 == 0:
                plan = {0: (Command.TAKEOFFYAW, [0.4, 1.5, 0.3]), 12: (Command.NOTIFY, [t]), 14: (Command.GOTO, [[1.2, 1.1, 0.6], 0.2, 1.0, False]),
                        40: (Command.GOTO, [[0.1, -0.1, 0.1], -0.1, 2.0, True]), 70: (Command.LANDVEL, [0.1, 0.3, False]),
                        90: (Command.STOP, [t]), 100: (Command.TAKEOFFVEL, [0.2, 0.4, True]), 130: (Command.LANDYAW, [0.0, 2.0, 0.0])}
            else:
                plan = {0: (Command.TAKEOFFVEL, [0.5, 0.4, False]), 25: (Command.NOTIFY, [t]), 26: (Command.LAND, [0.0, 2.0]),
                        40: (Command.GOTO, [[0.8, 0.7, 0.5], 0.0, 1.5, False]), 120: (Command.TAKEOFF, [0.6, 2.0]),
                        140: (Command.TAKEOFFVEL, [0.1, 0.0, True])}
            if k in plan:
                return plan[k]
            if i % 2 == 0 and 20 <= k < 38 or i % 2 == 1 and (95 <= k < 110 or 121 <= k < 125):
                return (Command.FULLSTATE, [p + [0.0, 0.0, 0.02], np.zeros(3), np.zeros(3), 0.1, np.zeros(3), t])
            return (Command.NONE, [])
        return act

    _commander_run(name, level, M, steps, seed, lambda obs, info: [script(i) for i in range(M)])


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--commander":
        commander_hardcoded_case("race_commander_hardcoded_gs", "getting_started", 2, 520, 8)
        commander_script_case("race_commander_script_gs", "getting_started", 2, 160, 9)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "--levels":
        # the remaining level configs: level0 (drone state randomised, nothing else: its own reset-draw layout), level1 on the
        # Bullet body (inertia + state randomisation and disturbances, nominal track), twogates (another track) with Physics.PYB_GND
        race_case("race_pyb_compete_level0", "level0", Physics.PYB, RaceMode.COMPETE, 2, 120, 11)
        race_case("race_pyb_compare_level1", "level1", Physics.PYB, RaceMode.COMPARE, 2, 120, 12)
        race_case("race_pybgnd_compare_twogates", "twogates", Physics.PYB_GND, RaceMode.COMPARE, 2, 120, 13)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "--fullstate":
        fullstate_case("race_fullstate_compare_gs", "getting_started", 2, 200, 6)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "--new":
        wrapper_case("race_wrappers_compare_gs", "getting_started", 2, 400, 5)
        return
    race_case("race_pyb_compete_gs", "getting_started", Physics.PYB, RaceMode.COMPETE, 2, 230, 1)
    race_case("race_pyb_compare_level3", "level3", Physics.PYB, RaceMode.COMPARE, 2, 150, 2)
    race_case("race_dyn_compete_level1", "level1", Physics.DYN, RaceMode.COMPETE, 2, 60, 3)
    race_case("race_pyb_all_compete_level3", "level3", Physics.PYB_GND_DRAG_DW, RaceMode.COMPETE, 2, 100, 4)


if __name__ == "__main__":
    main()
