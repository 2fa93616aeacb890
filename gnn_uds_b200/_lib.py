"""ctypes binding of ``libgnnuds.so`` (C ABI declared in ``include/gnnuds.h``).

There is no CPU fallback: if the shared library is missing or no CUDA device is
visible, the product path raises.  Build with ``python -m gnn_uds_b200.build``
(or ``__graft_entry__.build()``).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgnnuds.so")

ABI_VERSION = 2
CONV_NONE, CONV_GCN, CONV_GAT = 0, 1, 2
REC_CONV1D, REC_GRU, REC_LSTM = 1, 2, 3
ACT = {"linear": 0, None: 0, "relu": 1, "tanh": 2, "sigmoid": 3}
PREC_FP32, PREC_BF16_TC = 0, 1
MEM_HOST, MEM_DEVICE = 0, 1
ROLL_MPC, ROLL_MBRL, ROLL_NORM = 1, 2, 3


class Config(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "abi_version", "n_node", "n_edge", "n_in", "b_in", "e_in", "n_out", "e_out",
        "seq_in", "seq_out", "embed", "hidden", "kernel_size", "n_sp_layer", "n_tp_layer",
        "conv", "recurrent", "activation", "resnet", "if_flood", "edge_fusion", "act",
        "graph_base", "tide", "n_act")] + [("epsilon", C.c_float)] + [(n, C.c_int32) for n in (
            "precision", "device", "max_chunk", "workspace_mb")]


_pi, _pf = C.POINTER(C.c_int32), C.POINTER(C.c_float)


class Graph(C.Structure):
    _fields_ = [("row_ptr", _pi), ("col_idx", _pi), ("val", _pf), ("n", C.c_int32), ("nnz", C.c_int32),
                ("e_row_ptr", _pi), ("e_col_idx", _pi), ("e_val", _pf), ("m", C.c_int32), ("e_nnz", C.c_int32),
                ("edges", _pi), ("act_of_edge", _pi), ("act_out_of_node", _pi), ("act_in_of_node", _pi),
                ("hmin", _pf), ("hmax", _pf), ("is_outfall", _pf), ("ehmax", _pf),
                ("pump", _pf), ("offset", _pf), ("pump_in", _pf), ("pump_out", _pf), ("area", _pf)]


class GnnUdsError(RuntimeError):
    pass


_lib = None

# every symbol include/gnnuds.h declares (tests check the export list against this)
SYMBOLS = [
    "gnnuds_create", "gnnuds_destroy", "gnnuds_last_error", "gnnuds_set_graph", "gnnuds_weight_layout",
    "gnnuds_set_weights", "gnnuds_set_norm", "gnnuds_output_dims", "gnnuds_predict", "gnnuds_simulate",
    "gnnuds_rollout", "gnnuds_objective", "gnnuds_objective_static", "gnnuds_evaluate", "gnnuds_replay", "gnnuds_build_windows", "gnnuds_alloc_pinned", "gnnuds_free_pinned", "gnnuds_alloc_device",
    "gnnuds_free_device", "gnnuds_memcpy", "gnnuds_synchronize", "gnnuds_launch_count", "gnnuds_last_timing",
    "gnnuds_set_profiling", "gnnuds_measure_simt_tflops", "gnnuds_kernel_timing", "gnnuds_debug_stop_after", "gnnuds_debug_fetch", "gnnuds_debug_umma_selftest",
    "gnnuds_tc_operand_is_half", "gnnuds_set_post_proc_variant", "gnnuds_set_adj_action",
    "gnnuds_predict_vjp", "gnnuds_grad_settings", "gnnuds_replay_batch",
]


def load():
    """dlopen libgnnuds.so and set prototypes.  Raises if the extension is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GnnUdsError("%s not found: build the CUDA extension first (python -m gnn_uds_b200.build); "
                          "there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    P = C.POINTER
    lib.gnnuds_create.argtypes = [P(Config), P(vp)]
    lib.gnnuds_destroy.argtypes = [vp]
    lib.gnnuds_last_error.argtypes = [vp]
    lib.gnnuds_last_error.restype = C.c_char_p
    lib.gnnuds_set_graph.argtypes = [vp, P(Graph)]
    lib.gnnuds_weight_layout.argtypes = [vp, P(i32), P(i64)]
    lib.gnnuds_set_weights.argtypes = [vp, vp, i64, P(i64), P(i64), i32]
    lib.gnnuds_set_norm.argtypes = [vp, vp, i32, vp, i32, vp, i32, vp, i32]
    lib.gnnuds_output_dims.argtypes = [vp, P(i32), P(i32)]
    lib.gnnuds_predict.argtypes = [vp, vp, vp, vp, vp, i64, vp, vp, i32]
    lib.gnnuds_simulate.argtypes = [vp, vp, vp, i32, vp, vp, i64, vp, vp, i32]
    lib.gnnuds_rollout.argtypes = [vp, i32, vp, vp, vp, vp, i64, i32, vp, vp, i32]
    lib.gnnuds_objective.argtypes = [vp, vp, vp, i64, i32, vp, vp, vp, vp, vp, i32]
    lib.gnnuds_objective_static.argtypes = [vp, vp, vp, i64, i32, vp, vp, vp, vp, vp, i32]
    lib.gnnuds_alloc_pinned.argtypes = [P(vp), i64]
    lib.gnnuds_free_pinned.argtypes = [vp]
    lib.gnnuds_alloc_device.argtypes = [vp, P(vp), i64]
    lib.gnnuds_free_device.argtypes = [vp, vp]
    lib.gnnuds_memcpy.argtypes = [vp, vp, vp, i64, i32]
    lib.gnnuds_synchronize.argtypes = [vp]
    lib.gnnuds_launch_count.argtypes = [vp]
    lib.gnnuds_launch_count.restype = i64
    lib.gnnuds_last_timing.argtypes = [vp, P(C.c_float), P(C.c_float)]
    lib.gnnuds_set_profiling.argtypes = [vp, i32]
    lib.gnnuds_evaluate.argtypes = [vp, i32, vp, vp, vp, i32, vp, i32, i32, i64, i32, vp, vp, vp, vp, vp, i32, i32]
    lib.gnnuds_predict_vjp.argtypes = [vp, vp, vp, vp, vp, i64, vp, vp, vp, vp, vp, vp, vp, i32]
    lib.gnnuds_grad_settings.argtypes = [vp, i32, vp, vp, vp, vp, i64, i32, vp, vp, vp, vp, vp, vp, i32]
    lib.gnnuds_replay.argtypes = [vp, vp, vp, vp, vp, i64, vp, vp, i32]
    lib.gnnuds_replay_batch.argtypes = [vp, i32, P(vp), P(vp), P(vp), P(vp), P(i64), P(vp), P(vp)]
    lib.gnnuds_build_windows.argtypes = [vp, vp, vp, vp, vp, i64, i64, i64, vp, vp, vp, vp, i32]
    lib.gnnuds_measure_simt_tflops.argtypes = [vp, P(C.c_float)]
    lib.gnnuds_kernel_timing.argtypes = [vp, i32, P(C.c_float), P(i64)]
    lib.gnnuds_debug_stop_after.argtypes = [vp, i32]
    lib.gnnuds_set_post_proc_variant.argtypes = [vp, i32]
    lib.gnnuds_set_adj_action.argtypes = [vp, vp, vp]
    lib.gnnuds_debug_fetch.argtypes = [vp, i32, vp, i64, P(i64), P(i32)]
    lib.gnnuds_debug_umma_selftest.argtypes = [vp, i32, vp, i32, i32, vp, i32, i32, i32, i32, vp]
    for s in SYMBOLS:
        if s not in ("gnnuds_last_error", "gnnuds_launch_count"):
            getattr(lib, s).restype = i32
    _lib = lib
    return lib


def check(handle, rc):
    if rc != 0:
        msg = load().gnnuds_last_error(handle)
        raise GnnUdsError("libgnnuds error %d: %s" % (rc, msg.decode() if msg else "?"))
