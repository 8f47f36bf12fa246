"""ctypes binding of libdincae_b200.so -- one declaration per entry point of include/dincae_b200.h."""
import ctypes as C
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB_PATH = os.path.join(HERE, "libdincae_b200.so")
HEADER = os.path.join(ROOT, "include", "dincae_b200.h")

OK, EINVAL, EUNSUPPORTED, ECUDA, ENOMEM, ESIZE = range(6)
HOST, DEVICE, DEVICE_XIN = 0, 1, 2
UP_NEAREST, UP_BILINEAR = 0, 1
PREC_FP32, PREC_TF32, PREC_TF32X1 = 0, 1, 2
MAX_LEVELS, MAX_CHAINS = 8, 4
COMM_ID_BYTES = 128


class Config(C.Structure):
    _fields_ = [("W", C.c_int32), ("H", C.c_int32), ("ninput", C.c_int32), ("noutput", C.c_int32),
                ("nlevels", C.c_int32), ("nfilter", C.c_int32 * MAX_LEVELS), ("skip_mask", C.c_uint32),
                ("upsampling", C.c_int32), ("nchains", C.c_int32), ("loss_weights", C.c_float * MAX_CHAINS),
                ("gamma", C.c_float), ("truth_uncertain", C.c_int32), ("l1_beta", C.c_float), ("l2_beta", C.c_float),
                ("laplacian_penalty", C.c_float), ("laplacian_error_penalty", C.c_float), ("clip_grad", C.c_float),
                ("adam_beta1", C.c_float), ("adam_beta2", C.c_float), ("adam_eps", C.c_float),
                ("leaky_slope", C.c_float), ("precision", C.c_int32), ("max_batch", C.c_int32)]


class DincaeError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libdincae_b200 error %d: %s" % (code, msg))
        self.code = code


vp, i32, i64, u64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_float
pp = C.POINTER(C.c_void_p)

# name -> (restype, argtypes); pointers to caller data are void* so host (numpy) and device (int) both pass
SIGNATURES = {
    "dincae_ctx_create": (i32, [i32, pp]),
    "dincae_ctx_destroy": (i32, [vp]),
    "dincae_ctx_set_stream": (i32, [vp, vp]),
    "dincae_sync": (i32, [vp]),
    "dincae_last_error": (C.c_char_p, []),
    "dincae_version": (i32, []),
    "dincae_launch_count": (i64, [vp]),
    "dincae_profile_begin": (i32, [vp]),
    "dincae_profile_end": (i32, [vp, C.c_char_p, i64]),
    "dincae_interp": (i32, [vp, vp, i64, vp, i32, i32, vp, i32]),
    "dincae_interp_adjoint": (i32, [vp, vp, vp, i64, i32, i32, vp, i32]),
    "dincae_interp_adjoint_noalloc": (i32, [vp, vp, vp, i64, i32, i32, vp]),
    "dincae_interp_cells": (i32, [vp, vp, i64, vp, i32]),
    "dincae_model_create": (i32, [vp, C.POINTER(Config), pp]),
    "dincae_model_destroy": (i32, [vp]),
    "dincae_model_num_params": (i32, [vp, C.POINTER(i64)]),
    "dincae_model_get_params": (i32, [vp, vp, i64]),
    "dincae_model_set_params": (i32, [vp, vp, i64]),
    "dincae_model_init_glorot": (i32, [vp, u64]),
    "dincae_model_get_opt_state": (i32, [vp, vp, vp, i64, C.POINTER(i64)]),
    "dincae_model_set_opt_state": (i32, [vp, vp, vp, i64, i64]),
    "dincae_forward": (i32, [vp, vp, i32, vp, i32]),
    "dincae_loss": (i32, [vp, vp, vp, i32, i32, C.POINTER(C.c_double)]),
    "dincae_train_step": (i32, [vp, vp, vp, i32, f32, i32, C.POINTER(f32)]),
    "dincae_model_step_graph": (i32, [vp, i32, C.POINTER(i32)]),
    "dincae_vector2_loss": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, C.POINTER(C.c_double), vp]),
    "dincae_batch_upload": (i32, [vp, i32, vp, vp, i32]),
    "dincae_train_step_slot": (i32, [vp, i32, i32, f32, C.POINTER(f32)]),
    "dincae_grad_step_slot": (i32, [vp, i32, i32, vp]),
    "dincae_loss_points": (i32, [vp, vp, vp, vp, vp, i32, i32, C.POINTER(C.c_double)]),
    "dincae_train_step_points": (i32, [vp, vp, vp, vp, vp, i32, f32, i32, C.POINTER(f32)]),
    "dincae_grad_step": (i32, [vp, vp, vp, i32, vp, i32]),
    "dincae_grad_step_points": (i32, [vp, vp, vp, vp, vp, i32, i64, i32]),
    "dincae_grad_buffer": (i32, [vp, pp, C.POINTER(i64)]),
    "dincae_apply_step": (i32, [vp, f32, C.POINTER(f32)]),
    "dincae_comm_unique_id": (i32, [vp]),
    "dincae_comm_init_rank": (i32, [vp, i32, i32, vp]),
    "dincae_comm_destroy": (i32, [vp]),
    "dincae_comm_info": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]),
    "dincae_allreduce": (i32, [vp, vp, i64]),
    "dincae_allreduce_grads": (i32, [vp]),
    "dincae_train_step_dp": (i32, [vp, vp, vp, i32, vp, f32, i32, C.POINTER(f32)]),
    "dincae_train_step_dp_slot": (i32, [vp, i32, i32, vp, f32, C.POINTER(f32)]),
    "dincae_replica_divergence": (i32, [vp, C.POINTER(f32)]),
    "dincae_count_valid": (i32, [vp, vp, i32, i32, vp]),
    "dincae_gridded_create": (i32, [vp, vp, i32, i32, i32, i64, vp, vp, vp, i32, vp, vp, vp, vp, vp, vp, i32, i32, pp]),
    "dincae_gridded_destroy": (i32, [vp]),
    "dincae_gridded_set_error": (i32, [vp, vp, i32]),
    "dincae_gridded_get_mean": (i32, [vp, vp]),
    "dincae_gridded_assemble": (i32, [vp, vp, vp, u64, u64, i32, vp, vp, i32]),
    "dincae_train_step_gridded": (i32, [vp, vp, vp, vp, u64, u64, i32, f32, C.POINTER(f32)]),
    "dincae_train_step_dp_gridded": (i32, [vp, vp, vp, vp, u64, u64, i32, vp, f32, C.POINTER(f32)]),
    "dincae_gridded_valid_counts": (i32, [vp, i32, vp]),
    "dincae_reconstruct_range": (i32, [vp, vp, i64, i64, i32, f32, vp, vp, i32]),
    "dincae_points_create": (i32, [vp, i32, i32, i64, vp, vp, vp, vp, i32, vp, vp, vp, vp, i32, f32, f32, pp]),
    "dincae_points_destroy": (i32, [vp]),
    "dincae_posindices": (i32, [vp, vp, vp, i64, C.c_double, C.c_double, i32, C.c_double, C.c_double, i32, vp, vp, i32]),
    "dincae_points_assemble": (i32, [vp, vp, i32, i32, u64, u64, vp, vp, vp, i32]),
}

_lib = None


def header_symbols():
    """Every function name declared in include/*.h (the drop-in boundary)."""
    names = []
    inc = os.path.join(ROOT, "include")
    for fn in sorted(os.listdir(inc)):
        if fn.endswith(".h"):
            txt = open(os.path.join(inc, fn)).read()
            txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
            names += re.findall(r"\b(dincae_\w+)\s*\(", txt)
    return sorted(set(names))


def load():
    """dlopen the CUDA library; fails loudly when it is missing (there is no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("libdincae_b200.so is not built (%s); run `python __graft_entry__.py`. "
                          "dincae_b200 has no CPU or PyTorch fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    missing = [n for n in header_symbols() if not hasattr(lib, n)]
    if missing:
        raise ImportError("libdincae_b200.so lacks symbols declared in include/*.h: %s" % missing)
    for name, (res, args) in SIGNATURES.items():
        if hasattr(lib, name):
            f = getattr(lib, name)
            f.restype = res
            f.argtypes = args
    _lib = lib
    return lib


def check(code):
    if code != OK:
        raise DincaeError(code, load().dincae_last_error().decode("utf-8", "replace"))


def ptr(a):
    """void* of a numpy array, a raw device address (int), a torch tensor, or None."""
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    return C.c_void_p(a.ctypes.data)
