"""GPU: adrp_step_rows_static (include/adrp.h) -- the control step without a handle and without any allocation, its SoA
working state in a file-scope `static __device__` array of the library -- against the existing entry points
(adrp_set_state + adrp_step + adrp_get_state on a handle of the same config), through the C ABI.

No counterpart in the reference.  Collected last on purpose (pytest -x).
The new entry point launches the same kernel instances as adrp_step, so the comparison is exact (tolerance 0; the parity
tests' own tolerances are 1e-9 relative in fp64 and 1e-4 in fp32)."""
import ctypes as C

import numpy as np
import pytest

from gym_pybullet_adrp_b200 import _abi
from gym_pybullet_adrp_b200.utils.enums import Physics

pytestmark = pytest.mark.gpu
HOVER = float(np.sqrt(9.8 * 0.03454 / (4 * 3.16e-10)))


def _cfg(N, M, physics, precision):
    xyz = [[1.2 * (k % 3), 1.2 * (k // 3), 0.5 + 0.25 * (k % 4)] for k in range(M)]
    return _abi.make_config(num_envs=N, num_drones=M, physics=physics, pyb_freq=240, ctrl_freq=30,
                            action_type=_abi.ACT_RPM_RAW, env_kind=_abi.ENV_CTRL, precision=precision, initial_xyzs=xyz)


def _static_step(lib, cfg, state, action):
    """One call of the new entry point; every buffer is the caller's (torch), the call itself allocates nothing."""
    import torch
    N, M = cfg.num_envs, cfg.drones_per_env
    dt = state.dtype
    out = torch.empty_like(state)
    obs = torch.empty((N, M, 20), dtype=dt, device=state.device)
    rew = torch.empty(N, dtype=dt, device=state.device)
    te = torch.empty(N, dtype=torch.uint8, device=state.device)
    tr = torch.empty(N, dtype=torch.uint8, device=state.device)
    rc = lib.adrp_step_rows_static(C.byref(cfg), state.data_ptr(), action.data_ptr(), out.data_ptr(), obs.data_ptr(),
                                   rew.data_ptr(), te.data_ptr(), tr.data_ptr(),
                                   C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.adrp_last_error(None).decode()
    return out, obs, rew, te, tr


@pytest.mark.parametrize("precision", [_abi.F32, _abi.F64])
@pytest.mark.parametrize("physics,M", [(Physics.DYN_GND_DRAG_DW, 8), (Physics.PYB_GND_DRAG_DW, 3), (Physics.PYB, 1)])
def test_static_step_equals_handle_step_two_calls_in_a_row(physics, M, precision):
    import torch
    from gym_pybullet_adrp_b200 import _lib
    lib = _lib.load()
    N = 257
    rng = np.random.default_rng(100 * M + precision)
    cfg = _cfg(N, M, physics, precision)
    h = _lib.Handle(cfg)
    st0 = h.get_state().clone()
    st0[:, :, 0:3] += torch.as_tensor(rng.uniform(-0.05, 0.05, (N, M, 3)), dtype=h.dtype, device=h.device)
    st0[:, :, 7:10] = torch.as_tensor(rng.uniform(-0.3, 0.3, (N, M, 3)), dtype=h.dtype, device=h.device)
    acts = [torch.as_tensor(HOVER * (1 + 0.01 * rng.uniform(-1, 1, (N, M, 4))), dtype=h.dtype, device=h.device)
            for _ in range(2)]
    h.set_state(st0)
    state = st0
    for k in range(2):                                    # two calls in a row: the second starts from the first's output
        want_obs, want_rew, want_te, want_tr, _ = [t.clone() for t in h.step(acts[k])]
        want_state = h.get_state()
        state, obs, rew, te, tr = _static_step(lib, cfg, state, acts[k])
        torch.cuda.synchronize()
        assert torch.equal(obs, want_obs), (k, (obs - want_obs).abs().max().item())
        assert torch.equal(state, want_state), (k, (state - want_state).abs().max().item())
        assert torch.equal(rew, want_rew) and torch.equal(te, want_te) and torch.equal(tr, want_tr)
    h.close()


def test_static_step_is_independent_of_what_the_last_call_left_behind():
    """The array keeps the last call's contents: a small batch after a large one of another shape and precision gives
    what it gives on its own."""
    import torch
    from gym_pybullet_adrp_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(1)
    small = _cfg(5, 3, Physics.PYB_GND_DRAG_DW, _abi.F32)
    h = _lib.Handle(small)
    st = h.get_state().clone()
    a = torch.as_tensor(HOVER * (1 + 0.01 * rng.uniform(-1, 1, (5, 3, 4))), dtype=torch.float32, device=h.device)
    first = _static_step(lib, small, st, a)
    big = _cfg(4000, 8, Physics.DYN_GND_DRAG_DW, _abi.F64)
    hb = _lib.Handle(big)
    _static_step(lib, big, hb.get_state(), torch.full((4000, 8, 4), HOVER * 1.3, dtype=torch.float64, device=h.device))
    again = _static_step(lib, small, st, a)
    torch.cuda.synchronize()
    for x, y in zip(first, again):
        assert torch.equal(x, y)
    h.close(); hb.close()


def test_static_step_rejects_what_does_not_fit_and_what_it_does_not_support():
    import torch
    from gym_pybullet_adrp_b200 import _lib
    lib = _lib.load()
    dummy = torch.zeros(64, device="cuda")
    p = dummy.data_ptr()
    big = _cfg(131072, 8, Physics.DYN, _abi.F32)          # 2^20 padded drones x 80 B = 80 MiB of scratch
    rc = lib.adrp_step_rows_static(C.byref(big), p, p, p, p, p, p, p, None)
    msg = lib.adrp_last_error(None).decode()
    assert rc == -1 and "ADRP_STATIC_SCRATCH_BYTES" in msg and str(8 << 20) in msg, (rc, msg)
    hover = _abi.make_config(num_envs=4, num_drones=1, physics=Physics.PYB, pyb_freq=240, ctrl_freq=30,
                             action_type=_abi.ACT_RPM, env_kind=_abi.ENV_HOVER, precision=_abi.F32)
    rc = lib.adrp_step_rows_static(C.byref(hover), p, p, p, p, p, p, p, None)
    assert rc == -4 and "ADRP_ENV_CTRL" in lib.adrp_last_error(None).decode()
    ok = _cfg(4, 1, Physics.DYN, _abi.F32)
    assert lib.adrp_step_rows_static(C.byref(ok), None, p, p, p, p, p, p, None) == -1
