"""
GPU tests of the sampling semantics: shard invariance (the analogue of Pyrado/tests/test_sampling.py:594-646,
same rollouts for any worker count), auto-reset (min_steps) sampling, stop_on_done=False, ragged batch sizes,
fp32 vs fp64 builds on the full-size batch through size-independent properties.
"""
import numpy as np
import pytest
import torch

from test_gpu_engine import DEV, _t, default_distr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def E():
    import simurlacra_b200.engine as eng

    return eng


def _fnn(E, d, hid, seed=0, scale=0.5):
    from simurlacra_b200 import _lib as L

    rng = np.random.RandomState(seed)
    sizes = [d["do"], hid, hid, d["da"]]
    w = (rng.randn(sum(sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(3))) * scale).astype(np.float32)
    return E.make_policy(L.POL_FNN, _t(w, torch.float32), sizes=sizes), w, sizes


@pytest.mark.parametrize("n_envs", [1, 31, 33, 130, 1000])
def test_ragged_batches_and_shard_invariance(E, n_envs):
    """Any split of the global env range gives bit-identical rollouts (Philox keys use the global index), for
    batch sizes that are not multiples of the warp / vector width."""
    from simurlacra_b200 import _lib as L

    d = E.dims(L.QQ)
    cfg = E.make_cfg(L.QQ, 1 / 250.0, 60, dtype=torch.float32, act_norm=True)
    pol, _, _ = _fnn(E, d, 32)
    distr = E.make_distr_table(default_distr("qq-su"))
    full = E.rollout(cfg, pol, n_envs, 61, device=DEV, distr=distr, seed=3)
    cut = n_envs // 3
    parts = [E.rollout(cfg, pol, cut, 61, device=DEV, distr=distr, seed=3),
             E.rollout(cfg, pol, n_envs - cut, 61, device=DEV, distr=distr, seed=3, env_offset=cut)] if cut else []
    if parts:
        for k in ("ret", "length", "params"):
            assert torch.equal(full[k], torch.cat([p[k] for p in parts]))
        assert torch.equal(full["traj"], torch.cat([p["traj"] for p in parts], dim=1))
        assert torch.equal(full["row_flag"], torch.cat([p["row_flag"] for p in parts], dim=1))
    if n_envs > 1:
        # rollouts within one batch differ (float32 returns may collide by chance for a handful)
        assert len(torch.unique(full["ret"])) >= 0.99 * n_envs and len(torch.unique(full["params"], dim=0)) == n_envs
    # the same call twice is deterministic
    again = E.rollout(cfg, pol, n_envs, 61, device=DEV, distr=distr, seed=3)
    assert torch.equal(full["traj"], again["traj"])


def test_auto_reset_rows_are_consistent_episodes(E, orc):
    """min_steps-style sampling: every lane fills its row budget with consecutive episodes (episode index
    e, e+1, ...), each closed by a TERMINAL row; the pieces equal single-episode rollouts of the oracle."""
    from simurlacra_b200 import _lib as L

    kind = L.BOB  # ball-on-beam episodes end early (ball leaves the beam), so lanes see several episodes
    d = E.dims(kind)
    N, T = 64, 400
    cfg = E.make_cfg(kind, 0.01, 500, dtype=torch.float64)
    pol = E.make_policy(L.POL_IDLE)
    res = E.rollout(cfg, pol, N, T, device=DEV, seed=9, auto_reset=True)
    flags = res["row_flag"].cpu().numpy()
    st, rw = res["state"].cpu().numpy(), res["rew"].cpu().numpy()
    n_eps, n_steps = 0, 0
    for n in range(N):
        rows = flags[:, n]
        assert rows[-1] in (L.ROW_TERMINAL, L.ROW_TRUNCATED)
        start, ep = 0, 0
        for r in range(T):
            if rows[r] >= L.ROW_TERMINAL:
                length = r - start
                if rows[r] == L.ROW_TERMINAL:
                    assert rows[r - 1] == L.ROW_DONE
                    ref = orc.rollout_batch(kind, 0.01, 500, 1, 500, policy=orc.make_policy(orc.POL_IDLE), seed=9,
                                            env_offset=n, episode=ep)
                    assert int(ref["length"][0]) == length
                    assert np.abs(st[start: r + 1, n] - ref["state"][: length + 1, 0]).max() < 1e-9
                    np.testing.assert_allclose(rw[start:r, n], ref["rew"][:length, 0], rtol=1e-9, atol=1e-12)
                    n_eps += 1
                n_steps += length
                start, ep = r + 1, ep + 1
        assert ep >= 2
    stats = res["stats"].cpu().numpy()
    assert stats[L.STAT_EPISODES] == n_eps and stats[L.STAT_STEPS] == n_steps
    assert stats[L.STAT_FAILED] == n_eps  # every complete ball-on-beam episode here ended by leaving the state space


def test_stop_on_done_false_keeps_stepping(E, orc):
    from simurlacra_b200 import _lib as L

    kind = L.OMO
    cfg = E.make_cfg(kind, 0.02, 40, dtype=torch.float64)
    N, H = 32, 100
    rng = np.random.RandomState(0)
    acts = rng.uniform(-40, 40, size=(H, N, 1))
    res = E.rollout(cfg, E.make_policy(L.POL_ACTIONS), N, H + 1, device=DEV, actions=_t(acts, torch.float64), seed=1,
                    stop_on_done=False)
    ref = orc.rollout_batch(kind, 0.02, 40, N, H, policy=orc.make_policy(orc.POL_ACTIONS), actions=acts, seed=1,
                            stop_on_done=False)
    flags = res["row_flag"].cpu().numpy()
    np.testing.assert_array_equal(flags[:H] == L.ROW_DONE, ref["done"].astype(bool))
    assert (flags[H] == L.ROW_TRUNCATED).all()
    assert np.abs(res["state"].cpu().numpy() - ref["state"]).max() < 1e-9
    np.testing.assert_allclose(res["rew"][:H].cpu().numpy(), ref["rew"], rtol=1e-9, atol=1e-9)


def test_full_size_fp32_properties(E):
    """BASELINE config 2 at full size (2^20 envs, 600 steps, DR, fused FNN [64,64]) through size-independent
    properties: every lane takes exactly n_rows - 1 - episodes steps in auto-reset mode, statistics add up, no NaN
    status, determinism of the checksum across two launches, and a random subset equals its own re-run as a shard."""
    from simurlacra_b200 import _lib as L

    d = E.dims(L.QQ)
    N, T = 1 << 20, 601
    cfg = E.make_cfg(L.QQ, 1 / 250.0, 600, dtype=torch.float32, act_norm=True)
    pol, _, _ = _fnn(E, d, 64, scale=0.3)
    distr = E.make_distr_table(default_distr("qq-su"))
    a = E.rollout(cfg, pol, N, T, device=DEV, distr=distr, seed=0, auto_reset=True, record=False)
    b = E.rollout(cfg, pol, N, T, device=DEV, distr=distr, seed=0, auto_reset=True, record=False)
    sa, sb = a["stats"].cpu().numpy(), b["stats"].cpu().numpy()
    assert sa[L.STAT_STEPS] == sb[L.STAT_STEPS] and sa[L.STAT_EPISODES] == sb[L.STAT_EPISODES]
    assert torch.equal(a["ret"], b["ret"]) and torch.equal(a["length"], b["length"])
    assert int(a["status"].abs().sum()) == 0
    assert torch.isfinite(a["ret"]).all()
    # each complete episode costs one closing row; lanes never idle in auto-reset mode
    assert sa[L.STAT_STEPS] <= N * (T - 1) and sa[L.STAT_STEPS] >= N * (T - 1) - sa[L.STAT_EPISODES] - N
    lo = 12345
    sub = E.rollout(cfg, pol, 4096, T, device=DEV, distr=distr, seed=0, env_offset=lo, auto_reset=True, record=False)
    assert torch.equal(sub["ret"], a["ret"][lo: lo + 4096]) and torch.equal(sub["length"], a["length"][lo: lo + 4096])


def test_c4_records_at_scale_properties(E):
    """BASELINE config 4 (QBallBalancer + 20-parameter DR + fused FNN[16,16], full StepSequence records; 2^18 lanes x 500
    steps = 12 GB of records through the bulk-store path) through size-independent properties: the row flags account for every
    row, step rows sum to the statistics, the first episode's rewards sum to the lane's return, the observation of the
    ball balancer IS the state on every written row, closing rows carry no action / reward, and a shard re-run alone
    reproduces its records bit for bit."""
    from simurlacra_b200 import _lib as L
    from simurlacra_b200.domain_randomization.default_randomizers import create_default_randomizer_qbb
    from simurlacra_b200.environments.pysim.quanser_ball_balancer import QBallBalancerSim

    N, H = 1 << 18, 500
    cfg = E.make_cfg(L.QBB, 0.01, H, dtype=torch.float32, act_norm=True)
    sizes = [8, 16, 16, 2]
    w = _t(np.random.RandomState(1).randn(sum(sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(3))) * 0.3, torch.float32)
    pol = E.make_policy(L.POL_FNN, w, sizes=sizes)
    distr = create_default_randomizer_qbb().table(QBallBalancerSim._param_names)
    res = E.rollout(cfg, pol, N, H + 1, device=DEV, distr=distr, seed=2, auto_reset=True)
    flag, st = res["row_flag"], res["stats"].cpu().numpy()
    assert int(res["status"].abs().sum()) == 0
    stepped = (flag == L.ROW_STEP) | (flag == L.ROW_DONE)
    closing = (flag == L.ROW_TERMINAL) | (flag == L.ROW_TRUNCATED)
    assert int(stepped.sum()) == st[L.STAT_STEPS] and bool((stepped | closing).all())  # auto-reset: no idle rows
    assert int((flag == L.ROW_TERMINAL).sum()) == st[L.STAT_EPISODES]
    # first episode of every lane: rewards up to its length sum to the lane's return
    t = torch.arange(H + 1, device=DEV).view(-1, 1)
    first = t < res["length"].view(1, -1)
    ret = (res["rew"] * first).sum(dim=0)
    assert float((ret - res["ret"]).abs().max()) < 1e-3 * max(1.0, float(res["ret"].abs().max()))
    assert bool(stepped[first].all())
    assert torch.equal(res["obs"], res["state"])  # QBB observes its state (quanser_ball_balancer.py observe), closing rows included
    assert float(res["act"][closing].abs().max()) == 0.0 and float(res["rew"][closing].abs().max()) == 0.0
    lo, n = 98304, 2048
    sub = E.rollout(cfg, pol, n, H + 1, device=DEV, distr=distr, seed=2, env_offset=lo, auto_reset=True)
    assert torch.equal(sub["traj"], res["traj"][:, lo:lo + n]) and torch.equal(sub["row_flag"], flag[:, lo:lo + n])


def test_records_placed_inside_a_wider_buffer(E):
    """sre_rollout_io.traj_env_stride / traj_env_offset: a launch writes its lanes into columns [offset, offset + n) of a
    wider [T, M, W] buffer (the learner's buffer of a multi-GPU gather, sharding.RecordGather): same records as a launch
    into its own buffer, the other columns untouched; both store paths (bulk tiles for the fused FNN, the 128-bit loop)."""
    from simurlacra_b200 import _lib as L

    H, n, M, off = 30, 96, 256, 64
    for kind, sizes, do in ((L.QBB, [8, 16, 16, 2], 8), (L.QQ, [6, 10, 1], 6)):
        cfg = E.make_cfg(kind, 0.01, H, dtype=torch.float32, act_norm=True)
        w = _t((np.random.RandomState(2).randn(sum(sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(len(sizes) - 1))) * 0.3), torch.float32)
        pol = E.make_policy(L.POL_FNN, w, sizes=sizes)
        own = E.rollout(cfg, pol, n, H + 1, device=DEV, seed=4, auto_reset=True)
        width = own["traj"].shape[2]
        wide_t = torch.full((H + 1, M, width), -7.0, dtype=torch.float32, device=DEV)
        wide_f = torch.full((H + 1, M), 77, dtype=torch.uint8, device=DEV)
        res = E.rollout(cfg, pol, n, H + 1, device=DEV, seed=4, auto_reset=True, traj=wide_t, row_flag=wide_f, traj_env_offset=off)
        assert torch.equal(wide_t[:, off:off + n], own["traj"]) and torch.equal(wide_f[:, off:off + n], own["row_flag"])
        assert torch.equal(res["obs"], own["obs"]) and torch.equal(res["ret"], own["ret"])
        assert bool((wide_t[:, :off] == -7.0).all()) and bool((wide_t[:, off + n:] == -7.0).all())
        assert bool((wide_f[:, :off] == 77).all()) and bool((wide_f[:, off + n:] == 77).all())


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_gpu_record_gather_over_peer_memory():
    """sharding.RecordGather on 2 GPUs (torchrun, NCCL + symmetric memory): the learner's buffer filled by both ranks'
    rollout kernels is bit-identical to the NCCL gather of the local record buffers (tools/gather_bench.py exits 1 else)."""
    import os
    import os.path as osp
    import socket
    import subprocess
    import sys

    root = osp.dirname(osp.dirname(osp.abspath(__file__)))
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                        "127.0.0.1", "--master-port", str(port), osp.join(root, "tools", "gather_bench.py"), "--envs", "4096",
                        "--workload", "qbb", "--reps", "1"], capture_output=True, text=True, timeout=280, env=dict(os.environ))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert '"parity_peer_equals_nccl_gather": true' in r.stdout
