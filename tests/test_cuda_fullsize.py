"""GPU tests at BASELINE.json's FULL sizes (the workloads `bench.py` times), through size-independent properties:

* envs never interact, so a batch that repeats K template envs must reproduce, bit for bit and in every
  copy, what a K-env handle computes -- and the K-env handle is checked against the oracle at the fp32 tolerance;
* results do not depend on how the env range is cut: a handle for envs [a, b) of the global range (`env_offset`, the
  Philox subsequence is the GLOBAL env index) equals rows a..b of the full batch bit for bit -- resets, randomised
  tracks, disturbances, flags and in-kernel auto-resets included;
* the chunk-pipelined host-buffer call equals the device call at full size.

The oracle only sees the K-env / slice problems (seconds of CPU work); everything at full size is compared on the GPU.
"""
import numpy as np
import pytest

import bench
from gym_pybullet_adrp_b200 import _abi
from tests import golden_cases as gc
from tests.oracle_lib import OracleEnv

pytestmark = pytest.mark.gpu


def _handle(cfg):
    from gym_pybullet_adrp_b200 import _lib
    return _lib.Handle(cfg)


@pytest.mark.parametrize("wl", ["c5", "c4"])
def test_full_size_batch_is_K_independent_templates(wl):
    """C5 (131 072 envs x 8, DYN+GND+DRAG+DW) and C4 (32 768 x 8, PYB_GND_DRAG_DW): K = 64 template envs tiled over the
    whole batch."""
    import torch
    N, M = bench.WORKLOADS[wl][0], bench.WORKLOADS[wl][1]
    S = bench.WORKLOADS[wl][3] // bench.WORKLOADS[wl][4]
    K, steps = 64, 10
    full, small = _handle(bench.make_cfg(wl, N)), _handle(bench.make_cfg(wl, K))
    oracle = OracleEnv(bench.make_cfg(wl, K))
    st = bench.initial_state(wl, K, M, 11)
    acts = bench.make_actions(wl, K, M, small.A, 11, steps)
    full.set_state(np.tile(st, (N // K, 1, 1)))
    small.set_state(st)
    oracle.set_state(st.astype(np.float64))
    for k in range(steps):
        a = torch.as_tensor(acts[k], device=full.device)
        of, rf, tf, uf, ff = full.step(a.repeat(N // K, 1, 1))
        os_, rs, ts, us, fs = small.step(a)
        assert torch.equal(of.view(N // K, K, M, full.O), os_.unsqueeze(0).expand(N // K, -1, -1, -1)), (wl, k)
        assert torch.equal(ff.view(N // K, K, M), fs.unsqueeze(0).expand(N // K, -1, -1))
        assert torch.equal(tf.view(N // K, K), ts.unsqueeze(0).expand(N // K, -1))
        oc = oracle.step(acts[k].astype(np.float64), nthreads=4)[0]
        gc.assert_fp32_north_star(os_.cpu().numpy(), oc, "ctrl", (f"full_size_templates[{wl}]", k))   # 1e-4 m / 1e-4 rad
    assert full.launch_count >= steps and S >= 5
    full.close(); small.close(); oracle.close()


def test_full_size_hover_aviary_c2():
    """C2 (HoverAviary, 65 536 envs x 1, PYB, RPM actions, 72-float observation rows with the action buffer, auto-reset):
    actions with period K over the batch; rewards, flags and the action ring included."""
    import torch
    wl = "c2"
    N, K, steps = bench.WORKLOADS[wl][0], 128, 40                          # 40 steps: past the 0.5 s action-buffer depth
    full, small = _handle(bench.make_cfg(wl, N)), _handle(bench.make_cfg(wl, K))
    oracle = OracleEnv(bench.make_cfg(wl, K))
    full.reset(); small.reset(); oracle.reset()
    acts = bench.make_actions(wl, K, 1, small.A, 12, steps)
    for k in range(steps):
        a = torch.as_tensor(acts[k], device=full.device)
        of, rf, tf, uf, _ = full.step(a.repeat(N // K, 1, 1))
        os_, rs, ts, us, _ = small.step(a)
        assert torch.equal(of.view(N // K, K, 1, full.O), os_.unsqueeze(0).expand(N // K, -1, -1, -1)), k
        assert torch.equal(rf.view(N // K, K), rs.unsqueeze(0).expand(N // K, -1))
        assert torch.equal(tf.view(N // K, K), ts.unsqueeze(0).expand(N // K, -1))
        assert torch.equal(uf.view(N // K, K), us.unsqueeze(0).expand(N // K, -1))
        oc, rc, tc, uc, _ = oracle.step(acts[k].astype(np.float64), nthreads=4)
        gc.assert_fp32_north_star(os_.cpu().numpy(), oc, "kin", ("full_size_hover_c2", k))            # 1e-4 m / 1e-4 rad
        np.testing.assert_array_equal(us.cpu().numpy(), uc)
    full.close(); small.close(); oracle.close()


@pytest.mark.parametrize("wl,steps", [("c3", 50), ("c5m", 6)])
def test_full_size_race_is_cut_invariant(wl, steps):
    """C3 (level3 COMPETE, 16 384 envs x 2, every randomisation and the Philox disturbances on, auto-reset) over 2 s of
    flight -- eliminations and in-kernel resets happen -- and C5M (131 072 x 8 with Mellinger): a handle for a slice of
    the global env range reproduces the rows of the full batch bit for bit."""
    import torch
    N, M = bench.WORKLOADS[wl][0], bench.WORKLOADS[wl][1]
    a0, n_slice = (N // 2 + 37) // 1, 192
    full = _handle(bench.make_cfg(wl, N))
    part = _handle(bench.make_cfg(wl, n_slice, env_offset=a0))
    acts = bench.make_actions(wl, N, M, full.A, 13, 4)
    full.reset(); part.reset()
    assert torch.equal(full.obs[a0:a0 + n_slice], part.obs)                # randomised tracks / start poses of the slice
    resets = 0
    for k in range(steps):
        a = torch.as_tensor(acts[k % 4], device=full.device)
        of, rf, tf, uf, ff = full.step(a)
        op, rp, tp, up, fp = part.step(a[a0:a0 + n_slice])
        assert torch.equal(of[a0:a0 + n_slice], op), (wl, k)
        assert torch.equal(ff[a0:a0 + n_slice], fp) and torch.equal(tf[a0:a0 + n_slice], tp)
        assert torch.equal(rf[a0:a0 + n_slice], rp) and torch.equal(uf[a0:a0 + n_slice], up)
        resets += int(tf.sum().item())
        # terminated <=> every drone of the env is eliminated or finished (MultiRaceAviary.py:674-698)
        done = ((ff & (_abi.FLAG_ELIMINATED | _abi.FLAG_FINISHED)) != 0).all(dim=1)
        assert torch.equal(done, tf.bool())
    if wl == "c3":
        assert resets > 0                                                  # the horizon covers terminations + auto-resets
    assert torch.equal(full.get_state()[a0:a0 + n_slice], part.get_state())
    full.close(); part.close()


def test_full_size_host_path_equals_device_path():
    """adrp_step_host at C5's size (8 chunks over two streams, 17 MB in / 86 MB out per step) against adrp_step."""
    import torch
    wl = "c5"
    N, M = bench.WORKLOADS[wl][0], bench.WORKLOADS[wl][1]
    dev, host = _handle(bench.make_cfg(wl, N)), _handle(bench.make_cfg(wl, N))
    st = bench.initial_state(wl, N, M, 14)
    dev.set_state(st); host.set_state(st)
    acts = bench.make_actions(wl, N, M, dev.A, 14, 3)
    for k in range(3):
        od, rd, td, ud, fd = dev.step(acts[k])
        oh, rh, th, uh, fh = host.step_host(acts[k])
        assert torch.equal(od.cpu(), torch.from_numpy(oh)), k
        assert torch.equal(fd.cpu(), torch.from_numpy(fh)) and torch.equal(td.cpu(), torch.from_numpy(th))
    dev.close(); host.close()


@pytest.mark.parametrize("wl,K", [("c5", 8), ("c4", 10)])
def test_full_size_step_n_equals_single_steps(wl, K):
    """adrp_step_n at BASELINE's full batch sizes: K control steps in ONE launch against K single launches of a twin
    handle on the same 2^20-drone (C5) and 262 144-drone (C4) batches.  fp32 product build: the K-step kernels are separately
    compiled instances, so floats agree to rounding (2e-5, relative above 1, over these gentle flights).  (The RL aviaries'
    K-step path -- action ring, in-kernel auto-resets -- is compared with exact flags in tests/test_cuda_step_n.py at a few hundred
    envs and in fp64: at 65 536 envs some env always sits within rounding of a truncation threshold, where two fp32 instances may
    legitimately disagree on a reset.)"""
    import torch
    N, M = bench.WORKLOADS[wl][0], bench.WORKLOADS[wl][1]
    a, b = _handle(bench.make_cfg(wl, N)), _handle(bench.make_cfg(wl, N))
    if bench.WORKLOADS[wl][5] == _abi.ENV_CTRL:
        st = torch.as_tensor(bench.initial_state(wl, N, M, 3), device=a.device)
        a.set_state(st); b.set_state(st)
    pool = torch.as_tensor(bench.make_actions(wl, N, M, a.A, 3, 4), device=a.device)
    acts = torch.stack([pool[k % 4] for k in range(K)])
    del pool
    for k in range(K):
        oa, ra, ta, ua, fa = a.step(acts[k])
    b.step_n(acts)                                        # outputs overwritten in place: the last step's remain
    torch.cuda.synchronize()

    def close(x, y):
        return ((x - y).abs() / x.abs().clamp(min=1.0)).max().item()

    assert close(oa, b.obs) < 2e-5, (wl, close(oa, b.obs))
    assert close(ra, b.reward) < 1e-4
    assert torch.equal(ta, b.terminated) and torch.equal(ua, b.truncated) and torch.equal(fa, b.flags)
    assert close(a.get_state(), b.get_state()) < 2e-5
    # ... and the twins continue in lock step (counters, ring head and controller state were written back)
    oa2, _, _, ua2, _ = a.step(acts[0])
    ob2, _, _, ub2, _ = b.step(acts[0])
    assert close(oa2, ob2) < 2e-5 and torch.equal(ua2, ub2)
    a.close(); b.close()
