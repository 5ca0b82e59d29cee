"""ctypes binding of libsmi.so (the C ABI declared in include/smi.h).

This is the binding a maintainer of the reference would add (INTEGRATION.md shows the same
prototypes as a jax.ffi registration).  There is no CPU fallback: if the shared library is
missing or the device is not sm_100 every entry point raises.
"""

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsmi.so")

SMI_MAX_SITES = 16

MEDIAN_EXACT_MIDPOINT = 0
MEDIAN_JAX_FP32_INDEX = 1
PATH_AUTO, PATH_SMALL, PATH_TC = 0, 1, 2
MODEL_GAUSS, MODEL_BNN_REG, MODEL_BNN_CLS = 0, 1, 2
ACT_RELU, ACT_TANH = 0, 1
NOISE_GAMMA_PREC, NOISE_FIXED = 0, 1
GUIDE_NORMAL, GUIDE_DELTA = 0, 1
T_IDENTITY, T_EXP = 0, 1


class SmiSite(C.Structure):
    _fields_ = [("size", C.c_int32), ("transform", C.c_int32), ("lat", C.c_int32),
                ("loc", C.c_int32), ("rho", C.c_int32)]


class SmiModelDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("guide", C.c_int32), ("in_dim", C.c_int32),
                ("hidden", C.c_int32), ("out_dim", C.c_int32), ("act", C.c_int32),
                ("noise", C.c_int32), ("sigma", C.c_float), ("n_sites", C.c_int32),
                ("latent_dim", C.c_int32), ("particle_dim", C.c_int32),
                ("sites", SmiSite * SMI_MAX_SITES)]


class SmiError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libsmi error %d: %s" % (code, msg))
        self.code = code


_P = C.c_void_p
_I = C.c_int
_F = C.c_float
_D = C.c_double
_Z = C.c_size_t
_MD = C.POINTER(SmiModelDesc)

# name -> (restype, argtypes).  Every symbol include/smi.h declares appears here; the CPU test
# suite checks that the library exports each one.
PROTOTYPES = {
    "smi_version": (_I, []),
    "smi_last_error": (C.c_char_p, []),
    "smi_device_check": (_I, []),
    "smi_launch_count": (C.c_longlong, [_I]),
    "smi_profile": (_I, [_I]),
    "smi_profile_report": (_I, [C.c_char_p, _Z]),
    "smi_stage_create": (_I, [_Z, C.POINTER(C.c_void_p)]),
    "smi_stage_h2d": (_I, [_P, _P, _Z, _P, _P]),
    "smi_stage_h2dv": (_I, [_P, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), _I, _P, _P]),
    "smi_stage_destroy": (_I, [_P]),
    "smi_convert_u8_f32": (_I, [_P, _Z, _P, _P]),
    "smi_convert_i64_i32": (_I, [_P, _Z, _P, _P]),
    "smi_stein_force_workspace_bytes": (_Z, [_I, _I, _I]),
    "smi_stein_force_f32": (_I, [_P, _P, _I, _I, _F, _F, _I, _F, _I, _P, _P, _P, _P, _Z, _P]),
    "smi_sqdist_workspace_bytes": (_Z, [_I, _I, _I]),
    "smi_sqdist_f32": (_I, [_P, _I, _I, _I, _P, _P, _Z, _P]),
    "smi_median_bandwidth_workspace_bytes": (_Z, [_I]),
    "smi_median_bandwidth_f32": (_I, [_P, _I, _I, _P, _P, _P, _Z, _P]),
    "smi_median_bandwidth_noalloc_max_m": (_I, []),
    "smi_median_bandwidth_noalloc_f32": (_I, [_P, _I, _I, _P, _P, _P]),
    "smi_stein_rows_workspace_bytes": (_Z, [_I, _I, _I]),
    "smi_stein_rows_dist": (_I, [_P, _I, _I, _I, _I, _P, _Z, _P]),
    "smi_stein_rows_hist": (_I, [_I, _I, _I, _I, _I, _P, _P, _Z, _P]),
    "smi_stein_rows_scan": (_I, [_I, _I, _I, _I, _I, _I, _P, _P, _Z, _P]),
    "smi_stein_rows_tail": (_I, [_I, _I, _I, _I, _P, _P, _Z, _P]),
    "smi_stein_rows_bandwidth": (_I, [_I, _I, _I, _I, _I, _F, _F, _P, _P, _P, _P, _Z, _P]),
    "smi_stein_rows_force": (_I, [_P, _P, _I, _I, _I, _I, _F, _P, _P, _P, _Z, _P]),
    "smi_stein_sym_supported": (_I, [_I, _I, _I]),
    "smi_stein_sym_colsum": (_I, [_P, _I, _I, _I, _I, _P, _P, _Z, _P]),
    "smi_stein_sym_mean": (_I, [_P, _I, _I, _I, _I, _P, _Z, _P]),
    "smi_stein_sym_split": (_I, [_P, _I, _I, _I, _I, _I, _I, _P, _Z, _P]),
    "smi_stein_sym_gram": (_I, [_I, _I, _I, _I, _I, _I, _P, _Z, _P]),
    "smi_stein_sym_pack_t": (_I, [_I, _I, _I, _I, _I, _P, _P, _Z, _P]),
    "smi_stein_sym_unpack": (_I, [_P, _I, _I, _I, _I, _I, _P, _Z, _P]),
    "smi_adam_f32": (_I, [_P, _P, _P, _P, _F, _P, _I, _D, _D, _D, _D, _Z, _P]),
    "smi_adagrad_f32": (_I, [_P, _P, _P, _P, _F, _P, _I, _D, _D, _Z, _P]),
    "smi_elbo_grad_workspace_bytes": (_Z, [_MD, _I, _I, _I, _I]),
    "smi_elbo_grad_f32": (_I, [_MD, _P, _P, _P, _P, _P, _I, _I, _I, _F, _F, _I, _I, _P, _P, _P, _Z, _P]),
    "smi_elbo_grad_tau_f32": (_I, [_MD, _P, _P, _P, _P, _P, _I, _I, _I, _F, _P, _I, _I, _P, _P, _P, _Z, _P]),
    "smi_asvgd_tau_f32": (_I, [_P, _I, _I, _I, _I, _I, _P, _P]),
    "smi_prng_step_f32": (_I, [_MD, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P, _P]),
    "smi_prng_layout": (_I, [_I]),
    "smi_prng_randint": (_I, [_I]),
    "smi_prng_split": (_I, [_P, _I, _P, _P]),
    "smi_prng_uniform_f32": (_I, [_P, _Z, _F, _F, _P, _P]),
    "smi_prng_normal_f32": (_I, [_P, _Z, _P, _P]),
    "smi_prng_bits_u32": (_I, [_P, _Z, _P, _P]),
    "smi_loss_record_f32": (_I, [_P, _P, _P, _I, _P]),
    "smi_adam_gated_f32": (_I, [_P, _P, _P, _P, _F, _P, _D, _D, _D, _D, _Z, _P, _P]),
    "smi_adagrad_gated_f32": (_I, [_P, _P, _P, _P, _F, _D, _D, _Z, _P, _P]),
    "smi_step_commit_f32": (_I, [_P, _P, _P, _I, _P, _P, _P, _P, _F, _F, _I, _P]),
    "smi_predictive_reg_f32": (_I, [_MD, _P, _I, _P, _I, _P, _I, _P, _P, _P, _P, _P, _P, _P]),
    "smi_metric_workspace_bytes": (_Z, [_I, _I]),
    "smi_metric_rmse_f32": (_I, [_P, _P, _I, _I, _P, _P, _Z, _P]),
    "smi_predictive_latents_f32": (_I, [_MD, _P, _I, _P, _I, _P, _P, _P]),
    "smi_predictive_cls_layer1_workspace_bytes": (_Z, [_MD, _I, _I]),
    "smi_predictive_cls_layer1_f32": (_I, [_MD, _P, _I, _P, _I, _P, _P, _Z, _P]),
    "smi_predictive_cls_head_f32": (_I, [_MD, _P, _P, _P, _I, _I, _P, _P, _P]),
    "smi_metric_classification_f32": (_I, [_P, _P, _P, _I, _I, _I, _I, _P, _P]),
    "smi_metric_nll_f32": (_I, [_P, _I, _I, _P, _P, _P, _Z, _P]),
}

_lib = None


def load():
    """Load libsmi.so (once).  Raises if it has not been built -- there is no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "libsmi.so is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or python -m smi_experiments_b200.csrc.build).  smi_experiments_b200 has no CPU path.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def last_error():
    return load().smi_last_error().decode("utf-8", "replace")


def check(rc):
    if rc != 0:
        raise SmiError(rc, last_error())


def call(name, *args):
    """Call an int-returning entry point and raise SmiError on a non-zero status."""
    check(getattr(load(), name)(*args))
