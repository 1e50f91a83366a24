"""
ctypes binding of the C-ABI in ``include/sre.h`` (``simurlacra_b200/lib/libsre_b200.so``).

There is no CPU fallback: if the shared library is missing the import of any compute path raises, and every
compute call returns a CUDA error when no device is present.
"""
import ctypes as C
import os
import os.path as osp

_HERE = osp.dirname(osp.abspath(__file__))
LIB_PATH = os.environ.get("SRE_B200_LIB") or osp.join(_HERE, "lib", "libsre_b200.so")  # override: kernel experiments
CSRC_DIR = osp.join(_HERE, "csrc")

# ---- constants (include/sre.h) ----
QQ, QCP, QBB, BOB, OMO, PEND = 0, 1, 2, 3, 4, 5
F32, F64 = 0, 1
MAX_DS, MAX_DO, MAX_DA, MAX_DP, MAX_NH, MAX_DISTR, MAX_LAYERS, MAX_WIDTH, MAX_FEATS = 8, 8, 2, 20, 2, 32, 8, 128, 12
F_SIMPLE_DYN = 1
F_QCP_STAB = 2
F_DISC_ACT = 4
REW_QUADR, REW_EXP, REW_SCALED_EXP = 0, 1, 2
FR_ENABLED, FR_STATE_DEP, FR_TIME_DEP, FR_ALWAYS_POS, FR_ALWAYS_NEG = 1, 2, 4, 8, 16
POL_ACTIONS, POL_FNN, POL_LINEAR, POL_IDLE = 0, 1, 2, 3
PREC_FP32, PREC_TF32X3, PREC_TF32 = 0, 1, 2
STEP_YIELDED_FINAL, STEP_COUNT_MASK = 1 << 30, 0x3FFFFFFF
ACT_NOISE_NONE, ACT_NOISE_NORMAL, ACT_NOISE_UNIFORM = 0, 1, 2
ENV_ACT_NOISE_NONE, ENV_ACT_NOISE_OUTER, ENV_ACT_NOISE_INNER = 0, 1, 2
FEATS = dict(const=0, identity=1, sign=2, abs=3, squared=4, cubic=5, sin=6, cos=7, sinsin=8, sincos=9, sig=10, bell=11,
             mult=12, atan2=13, rf=14, rbf=15)
MAX_FEAT_PARAMS = 1024
DISTR_NORMAL, DISTR_UNIFORM, DISTR_BERNOULLI = 0, 1, 2
ROW_STEP, ROW_DONE, ROW_TERMINAL, ROW_TRUNCATED, ROW_IDLE = 0, 1, 2, 3, 4
ST_NAN_STATE, ST_NAN_ACT, ST_NAN_OBS = 1, 2, 4
STAT_EPISODES, STAT_SUM_RET, STAT_SUM_RET2, STAT_SUM_LEN, STAT_STEPS, STAT_FAILED, STAT_MIN_RET, STAT_MAX_RET = range(8)
N_STATS = 8


class Task(C.Structure):
    _fields_ = [
        ("state_des", C.c_double * MAX_DS),
        ("Q", C.c_double * (MAX_DS * MAX_DS)),
        ("R", C.c_double * (MAX_DA * MAX_DA)),
        ("radial_mod", C.c_double * MAX_DS),
        ("min_rew", C.c_double),
        ("final_factor", C.c_double),
        ("rew_kind", C.c_int32),
        ("radial_mask", C.c_uint32),
        ("final_flags", C.c_int32),
        ("reserved", C.c_int32),
    ]


MAX_DELAY, MAX_DOWN, MAX_VF, MAX_AUG = 8, 8, 4, 8


class StepWrappers(C.Structure):
    """sre_step_wrappers: the stateful step-path wrappers the persistent rollout kernel runs per lane."""

    _fields_ = [
        ("act_delay", C.c_int32),
        ("act_delay_inner", C.c_int32),
        ("delay_first", C.c_int32),
        ("downsample", C.c_int32),
        ("obs_keep_mask", C.c_int32),
        ("vf_n", C.c_int32),
        ("vf_pos", C.c_int32 * MAX_VF),
        ("vf_vel", C.c_int32 * MAX_VF),
        ("vf_b0", C.c_double),
        ("vf_b1", C.c_double),
        ("vf_a1", C.c_double),
        ("vf_zi", C.c_double),
        ("n_aug", C.c_int32),
        ("aug_fixed", C.c_int32),
        ("aug_param", C.c_int32 * MAX_AUG),
        ("aug_nominal", C.c_double * MAX_AUG),
    ]


class EnvCfg(C.Structure):
    _fields_ = [
        ("kind", C.c_int32),
        ("dtype", C.c_int32),
        ("flags", C.c_int32),
        ("wild_init", C.c_int32),
        ("dt", C.c_double),
        ("max_steps", C.c_int64),
        ("task", Task),
        ("base_params", C.c_double * MAX_DP),
        ("act_norm", C.c_int32),
        ("obs_noise", C.c_int32),
        ("noise_std", C.c_double * MAX_DO),
        ("noise_mean", C.c_double * MAX_DO),
        ("has_init_box", C.c_int32),
        ("reserved", C.c_int32),
        ("init_box_lo", C.c_double * MAX_DS),
        ("init_box_up", C.c_double * MAX_DS),
        ("obs_norm", C.c_int32),
        ("env_act_noise", C.c_int32),
        ("obs_lb", C.c_double * MAX_DO),
        ("obs_ub", C.c_double * MAX_DO),
        ("env_act_noise_std", C.c_double * MAX_DA),
        ("env_act_noise_mean", C.c_double * MAX_DA),
        ("wrap", StepWrappers),
    ]


class Distr(C.Structure):
    _fields_ = [
        ("param_index", C.c_int32),
        ("kind", C.c_int32),
        ("a", C.c_float),
        ("b", C.c_float),
        ("c", C.c_float),
        ("clip_lo", C.c_float),
        ("clip_up", C.c_float),
        ("roundint", C.c_int32),
    ]


class Policy(C.Structure):
    _fields_ = [
        ("kind", C.c_int32),
        ("n_layers", C.c_int32),
        ("sizes", C.c_int32 * (MAX_LAYERS + 1)),
        ("n_feats", C.c_int32),
        ("feats", C.c_int32 * MAX_FEATS),
        ("hidden_nonlin", C.c_int32),
        ("output_nonlin", C.c_int32),
        ("cartpole_feats", C.c_int32),
        ("precision", C.c_int32),
        ("weights", C.c_void_p),
        ("n_weights", C.c_int64),
        ("act_noise", C.c_int32),
        ("reserved", C.c_int32),
        ("act_noise_scale", C.c_float * MAX_DA),
        ("act_noise_mean", C.c_float * MAX_DA),
        ("group_size", C.c_int64),
        ("n_groups", C.c_int64),
        ("feat_params", C.c_void_p),
        ("n_feat_params", C.c_int32),
        ("feat_arg", C.c_int32 * MAX_FEATS),
        ("weights_host", C.c_void_p),
    ]


class RolloutIO(C.Structure):
    _fields_ = [
        ("n_envs", C.c_int64),
        ("n_rows", C.c_int64),
        ("seed", C.c_uint64),
        ("env_offset", C.c_uint64),
        ("episode", C.c_uint32),
        ("stop_on_done", C.c_int32),
        ("auto_reset", C.c_int32),
        ("n_distr", C.c_int32),
        ("distr", C.c_void_p),
        ("params", C.c_void_p),
        ("init", C.c_void_p),
        ("n_init_cols", C.c_int32),
        ("reserved", C.c_int32),
        ("actions", C.c_void_p),
        ("traj", C.c_void_p),
        ("row_flag", C.c_void_p),
        ("ret", C.c_void_p),
        ("length", C.c_void_p),
        ("params_out", C.c_void_p),
        ("stats", C.c_void_p),
        ("status", C.c_void_p),
        ("traj_env_stride", C.c_int64),
        ("traj_env_offset", C.c_int64),
    ]


# every symbol include/sre.h declares: name -> (restype, argtypes)
_VP, _I, _I64, _U64, _U32, _D = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_uint32, C.c_double
_PI = C.POINTER(C.c_int)
SIGNATURES = {
    "sre_abi_version": (_I, []),
    "sre_last_error": (C.c_char_p, []),
    "sre_device_count": (_I, [_PI]),
    "sre_env_dims": (_I, [_I, _PI, _PI, _PI, _PI, _PI, _PI]),
    "sre_nominal_params": (_I, [_I, _I, _VP]),
    "sre_default_task": (_I, [_I, C.POINTER(Task)]),
    "sre_record_width": (_I, [_I]),
    "sre_record_layout": (_I, [_I, _PI, _PI, _PI, _PI, _PI, _PI]),
    "sre_stats_init": (_I, [_VP, _VP]),
    "sre_derive": (_I, [C.POINTER(EnvCfg), _VP, _I64, _VP, _VP, _VP, _VP, _VP, _VP]),
    "sre_sample_params": (_I, [C.POINTER(EnvCfg), _VP, _I, _U64, _U64, _U32, _VP, _I64, _VP]),
    "sre_sample_init": (_I, [C.POINTER(EnvCfg), _VP, _U64, _U64, _U32, _VP, _I64, _VP]),
    "sre_reset": (_I, [C.POINTER(EnvCfg), _VP, _VP, _I, _VP, _VP, _VP, _VP, _VP, _U64, _U64, _U32, _I64, _VP]),
    "sre_step": (_I, [C.POINTER(EnvCfg), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _U64, _U64, _U32, _I64, _VP]),
    "sre_observe": (_I, [C.POINTER(EnvCfg), _VP, _VP, _I64, _VP]),
    "sre_step_rew": (_I, [C.POINTER(EnvCfg), _VP, _VP, _VP, _VP, _I64, _VP]),
    "sre_task_step_rew": (_I, [C.POINTER(Task), _I, _I, _I, _D, _VP, _VP, _VP, _I64, _VP]),
    "sre_policy_forward": (_I, [C.POINTER(Policy), _VP, _I, _VP, _I, _I64, _VP]),
    "sre_rollout": (_I, [C.POINTER(EnvCfg), C.POINTER(Policy), C.POINTER(RolloutIO), _VP]),
    "sre_last_launch_count": (_I, []),
    "sre_returns": (_I, [_I, _VP, _I64, _I64, _VP, _VP, _D, _D, _VP, _VP, _I64, _I64, _VP]),
    "sre_qcp_diff_forward": (_I, [C.POINTER(EnvCfg), _VP, _VP, _VP, _VP, _I64, _I64, _VP, _VP, _VP, _VP, _VP]),
    "sre_qcp_diff_backward": (_I, [C.POINTER(EnvCfg), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _I64, _I64, _VP, _VP, _VP, _VP]),
    "sre_qcp_diff_jacobian": (_I, [C.POINTER(EnvCfg), _VP, _VP, _VP, _VP, _I64, _VP, _VP]),
    "sre_qcp_diff_grad_static": (_I, [C.POINTER(EnvCfg), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _I64, _I64, _VP, _VP, _VP, _VP,
                                      _VP, _VP]),
    "sre_bench_fma": (_I, [_I, _I64, _I, _I, _VP, C.POINTER(C.c_double), _VP]),
    "sre_bench_hbm_write": (_I, [_VP, _I64, _VP]),
}

_lib = None


class EngineError(RuntimeError):
    """A C-ABI call returned non-zero."""


def lib():
    """Load the shared library (once).  Fails loudly when it has not been built: there is no fallback."""
    global _lib
    if _lib is None:
        if not osp.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build the CUDA extension first (python -c 'import __graft_entry__ as g; "
                f"g.build()' or make -C {CSRC_DIR}). simurlacra_b200 has no CPU fallback."
            )
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(L, name)  # AttributeError if the library does not export a declared symbol
            f.restype = res
            f.argtypes = args
        if L.sre_abi_version() != 1:
            raise ImportError("libsre_b200.so ABI version mismatch")
        _lib = L
    return _lib


def check(code: int, what: str = ""):
    if code != 0:
        msg = lib().sre_last_error().decode()
        raise EngineError(f"{what} failed with code {code}: {msg}")
