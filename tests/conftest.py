"""
Test configuration.  `-m "not gpu"` covers the oracle against the golden vectors, the host-side logic and the
C-ABI export list; `-m gpu` holds the parity tests proper (they call the CUDA engine through the C-ABI).
"""
import os
import os.path as osp
import sys

import numpy as np
import pytest

ROOT = osp.dirname(osp.dirname(osp.abspath(__file__)))
for p in (ROOT, osp.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = osp.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden(name):
    return np.load(osp.join(GOLDEN, name + ".npz"), allow_pickle=True)


@pytest.fixture(scope="session")
def orc():
    import oracle as O

    O.build()
    return O


ENV_FILES = [
    # (golden file, kind name, ctor kwargs of the oracle env)
    ("traj_qq-su", "qq-su", {}),
    ("traj_qcp-su", "qcp-su", dict(long_pole=True)),
    ("traj_qcp-su_simple", "qcp-su", dict(long_pole=True, simple_dynamics=True)),
    ("traj_qbb", "qbb", {}),
    ("traj_qbb_simple", "qbb", dict(simple_dynamics=True)),
    ("traj_bob", "bob", {}),
    ("traj_omo", "omo", {}),
    ("traj_pend", "pend", {}),
    # stabilisation variants (same dynamics, other init / state spaces and tasks)
    ("traj_qq-st", "qq-su", dict(stab="qq-st")),
    ("traj_qcp-st", "qcp-su", dict(long_pole=True, stab="qcp-st")),
    ("traj_qcp-st_simple", "qcp-su", dict(long_pole=True, simple_dynamics=True, stab="qcp-st")),
    ("traj_bob-d", "bob", dict(disc_act=True)),  # BallOnBeamDiscSim: actions projected onto {-max, 0, +max}
    # non-chaotic regime of the swing-up pendulums: |a| <= 0.5 V about the hanging equilibrium, 600 steps (north_star's
    # fp32 bar -- 1e-4 over the FULL horizon, bit-exact lengths -- is asserted on these)
    ("traj_qq-su_calm", "qq-su", {}),
    ("traj_qcp-su_calm", "qcp-su", dict(long_pole=True)),
    ("traj_qcp-su_calm_simple", "qcp-su", dict(long_pole=True, simple_dynamics=True)),
]

# closed-loop FNN [do, 64, 64, da] goldens (tests/golden/fnn64_closed_loop_envs.npz): tag -> (kind name, oracle kwargs)
FNN64_CASES = {"qq": ("qq-su", {}), "qq_norm": ("qq-su", {}), "qcp": ("qcp-su", dict(long_pole=True)), "qbb": ("qbb", {}),
               "pend": ("pend", {})}

# task_args defaults of the stabilisation envs: quanser_qube.py:223-229 (ExpQuadrErrRewFcn), quanser_cartpole.py:491-500
# (QuadrErrRewFcn inside FinalRewTask(state_dependent, time_dependent)); RadiallySymmDesStateTask idcs=[1] in both
STAB_TASKS = {
    "qq-st": dict(state_des=[0.0, np.pi, 0.0, 0.0], Q=np.diag([3.0, 4.0, 2.0, 2.0]), R=np.diag([5e-2]), rew="exp", final=False),
    "qcp-st": dict(state_des=[0.0, np.pi, 0.0, 0.0], Q=np.diag([5e-0, 1e1, 1e-2, 1e-2]), R=np.diag([1e-3]), rew="quadr",
                   final=True),
}

OMO_TASK = dict(state_des=[0.5, 0.0], Q=np.diag([1e1, 1e-2]), R=np.diag([1e-6]))
