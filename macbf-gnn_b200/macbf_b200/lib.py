"""ctypes binding of libmacbf_b200.so (C ABI declared in include/macbf_b200.h).

There is no CPU or eager-PyTorch fallback: if the shared library is missing, or a tensor is not
on a CUDA device, the call raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MACBF_B200_LIB", os.path.join(os.path.dirname(_HERE), "libmacbf_b200.so"))

c_f32p = ctypes.c_void_p
c_i32p = ctypes.c_void_p
c_u8p = ctypes.c_void_p
c_void = ctypes.c_void_p
I, D, SZ = ctypes.c_int, ctypes.c_double, ctypes.c_size_t

# name -> (restype, argtypes); mirrors include/macbf_b200.h one to one
SIGNATURES = {
    "macbf_last_error": (ctypes.c_char_p, []),
    "macbf_version": (I, []),
    "macbf_launch_count": (ctypes.c_longlong, []),
    "macbf_timing_begin": (I, [c_void]),
    "macbf_timing_end": (I, [I, ctypes.c_char_p, I, ctypes.c_void_p]),
    "macbf_neighbor_count": (I, [I, I, I]),
    "macbf_neighbor_topk": (I, [c_f32p, c_f32p, I, I, I, I, c_i32p, c_void]),
    "macbf_neighbor_workspace_bytes": (SZ, [I, I, I, I]),
    "macbf_neighbor_topk_ws": (I, [c_f32p, c_f32p, I, I, I, I, c_i32p, c_void, SZ, c_void]),
    "macbf_neighbor_topk_part": (I, [c_f32p, c_f32p, I, I, I, I, I, I, c_i32p, c_void, SZ, c_void]),
    "macbf_dense_topk": (I, [c_f32p, I, I, I, I, c_i32p, c_void]),
    "macbf_cbf_forward": (I, [c_f32p, c_f32p, c_i32p, c_f32p, I, I, I, I, I, D, D, c_f32p, c_void]),
    "macbf_controller_forward": (I, [c_f32p, c_f32p, c_f32p, c_i32p, c_f32p, I, I, I, I, I, D, D, c_f32p, c_f32p,
                                     c_f32p, c_void]),
    "macbf_ttc_mask": (I, [c_f32p, c_f32p, c_i32p, I, I, I, I, D, D, c_u8p, c_void]),
    "macbf_loss_scratch_bytes": (SZ, [I, I, I]),
    "macbf_loss_forward": (I, [c_f32p, c_f32p, c_f32p, c_i32p, c_f32p, c_f32p, c_f32p, I, I, I, I, D, D, D, D, c_u8p,
                               c_f32p, c_void, SZ, c_void]),
    "macbf_loss_forward_noalloc": (I, [c_f32p, c_f32p, c_f32p, c_i32p, c_f32p, c_f32p, c_f32p, I, I, I, I, D, D, D, D,
                                       c_u8p, c_f32p, c_void]),
    "macbf_loss_backward": (I, [c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, c_u8p, c_f32p, c_f32p, I, I, I, D, D, c_f32p,
                                c_f32p, c_f32p, c_f32p, c_void]),
    "macbf_cbf_backward_scratch_bytes": (SZ, [I, I, I]),
    "macbf_cbf_backward": (I, [c_f32p, c_f32p, c_i32p, c_f32p, c_f32p, I, I, I, I, D, D, c_f32p, c_f32p, c_void, SZ,
                               c_void]),
    "macbf_controller_backward_scratch_bytes": (SZ, [I, I, I]),
    "macbf_controller_backward": (I, [c_f32p, c_f32p, c_f32p, c_i32p, c_f32p, c_f32p, c_f32p, I, I, I, I, D, D,
                                      c_f32p, c_f32p, c_f32p, c_void, SZ, c_void]),
    "macbf_scenario_workspace_bytes": (SZ, [I, I, I]),
    "macbf_generate_scenario": (I, [ctypes.c_ulonglong, I, I, I, D, c_f32p, c_f32p, c_i32p, c_void, SZ, c_void]),
    "macbf_ttc_dense": (I, [c_f32p, I, I, D, D, c_i32p, c_u8p, c_void]),
    "macbf_generate_obstacles": (I, [c_i32p, c_void, c_void, I, I, c_void, c_void]),
    "macbf_rollout_workspace_bytes": (SZ, [I, I, I, I]),
    "macbf_rollout_forward": (I, [c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, I, I, I, I, I, I, D, D, D, c_f32p, c_f32p, c_i32p,
                                  c_void, SZ, c_void]),
    "macbf_rollout_backward_scratch_bytes": (SZ, [I, I, I, I]),
    "macbf_rollout_backward": (I, [c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, I, I, I, I, I, D, D, c_f32p, c_f32p, c_f32p,
                                   c_void, SZ, c_void]),
    "macbf_step_workspace_bytes": (SZ, [I, I, I, I]),
    "macbf_step_workspace_winners_offset": (SZ, [I, I, I, I]),
    "macbf_controller_backward_saved": (I, [c_f32p, c_f32p, c_f32p, c_i32p, c_f32p, c_f32p, c_f32p, I, I, I, I, D, D, c_f32p,
                                            c_u8p, c_f32p, c_f32p, c_f32p, c_void, SZ, c_void]),
    "macbf_step_forward_rows_a": (I, [c_f32p, c_f32p, c_f32p, c_f32p, I, I, I, I, D, D, I, I, c_i32p, c_f32p, c_f32p, c_void, SZ,
                                      c_void]),
    "macbf_step_forward_rows_b": (I, [c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, I, I, I, I, D, D, D, D, D, I, I, c_i32p, c_i32p,
                                      c_f32p, c_f32p, c_f32p, c_u8p, c_f32p, c_void, SZ, c_void]),
    "macbf_step_forward_rows_c": (I, [c_f32p, I, I, c_f32p, c_void]),
    "macbf_cbf_record_bytes": (SZ, [I, I, I]),
    "macbf_cbf_backward_record_scratch_bytes": (SZ, []),
    "macbf_cbf_forward_record": (I, [c_f32p, c_f32p, c_i32p, c_f32p, I, I, I, I, D, D, c_f32p, c_void, SZ, c_void]),
    "macbf_cbf_backward_record": (I, [c_i32p, c_f32p, c_f32p, I, I, I, c_void, SZ, c_f32p, c_f32p, c_void, SZ, c_void]),
    "macbf_step_forward_train": (I, [c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, I, I, I, I, I, D, D, D, D, D, I, c_i32p, c_i32p,
                                     c_f32p, c_f32p, c_f32p, c_f32p, c_u8p, c_f32p, c_void, SZ, c_void, c_void, SZ, c_void]),
    "macbf_step_forward": (I, [c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, I, I, I, I, I, D, D, D, D, D, c_i32p, c_i32p,
                               c_f32p, c_f32p, c_f32p, c_f32p, c_u8p, c_f32p, c_void, SZ, c_void]),
    "macbf_step_forward_ex": (I, [c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, I, I, I, I, I, D, D, D, D, D, I, c_i32p, c_i32p,
                                  c_f32p, c_f32p, c_f32p, c_f32p, c_u8p, c_f32p, c_void, SZ, c_void]),
}

_lib = None


class MacbfError(RuntimeError):
    pass


def load():
    """Load (once) and return the ctypes handle.  Raises if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found - build it with macbf-gnn_b200/build.sh (or __graft_entry__.build()); "
            "there is no CPU fallback for this path")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().macbf_last_error()
        raise MacbfError(f"{what} failed (rc={rc}): {msg.decode() if msg else '?'}")
