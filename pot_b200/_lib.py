"""ctypes binding of libb2ot.so (the C ABI declared in include/b2ot.h).

The library is the only compute path of this package: if it is missing the import
fails loudly (there is no CPU or PyTorch fallback).
"""

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# B2OT_LIB: load another build of the same library (A/B measurements of kernel variants); default = the in-tree build
LIB_PATH = os.environ.get("B2OT_LIB") or os.path.join(_HERE, "libb2ot.so")

F32, F64 = 0, 1
MODE_KNOPP, MODE_LOG = 0, 1
ST_RUNNING, ST_CONVERGED, ST_NUMERR, ST_MAXITER = 0, 1, 2, 3
VARIANT_AUTO, VARIANT_FUSED, VARIANT_SPLIT, VARIANT_RESIDENT = 0, 1, 2, 3
OP_SUM, OP_MAX = 0, 1
BATCH_FORCE_EXACT = 1
VALUE_LINEAR, VALUE_ENTROPY, VALUE_KL, VALUE_KMULQ = 1, 2, 4, 8
MULTI_STATS, MULTI_KNOPP, MULTI_UNBALANCED = 0, 1, 2

ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_void_p)

# name -> (restype, argtypes); every symbol include/b2ot.h declares
PROTOTYPES = {
    "b2ot_version": (C.c_int, []),
    "b2ot_build_info": (C.c_char_p, []),
    "b2ot_last_error": (C.c_char_p, []),
    "b2ot_device_info": (C.c_int, [C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "b2ot_launch_count": (C.c_ulonglong, []),
    "b2ot_kexp": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_double, C.c_int,
                            C.c_void_p]),
    "b2ot_negscale": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_double,
                                C.c_int, C.c_void_p]),
    "b2ot_dense_workspace_bytes": (C.c_size_t, [C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int]),
    "b2ot_dense_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_int64,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                    C.c_int, C.c_int, C.c_double, C.c_int, C.c_void_p]),
    "b2ot_dense_set_unbalanced": (C.c_int, [C.c_void_p, C.c_double, C.c_double]),
    "b2ot_dense_set_stabilized": (C.c_int, [C.c_void_p, C.c_double, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "b2ot_dense_absorb_count": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "b2ot_dense_set_allreduce": (C.c_int, [C.c_void_p, ALLREDUCE_FN, C.c_void_p]),
    "b2ot_dense_init": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2ot_dense_iterate": (C.c_int, [C.c_void_p, C.c_int]),
    "b2ot_dense_profile": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "b2ot_dense_poll": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "b2ot_dense_solve": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int),
                                   C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "b2ot_dense_export": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_double), C.c_int,
                                    C.POINTER(C.c_int)]),
    "b2ot_dense_info": (C.c_int, [C.c_void_p] + [C.POINTER(C.c_int)] * 5),
    "b2ot_dense_destroy": (C.c_int, [C.c_void_p]),
    "b2ot_dense_plan": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                  C.c_int64, C.c_int64, C.c_int64, C.c_void_p]),
    "b2ot_dense_loss_scratch_bytes": (C.c_size_t, [C.c_int64, C.c_int64]),
    "b2ot_dense_loss": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                  C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2ot_dense_plan_value_scratch_bytes": (C.c_size_t, [C.c_int64, C.c_int64]),
    "b2ot_dense_plan_value": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                        C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_void_p,
                                        C.c_void_p, C.c_void_p]),
    "b2ot_colsum_scratch_bytes": (C.c_size_t, [C.c_int64, C.c_int64]),
    "b2ot_colsum": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                              C.c_void_p]),
    "b2ot_multi_scratch_bytes": (C.c_size_t, [C.c_int64, C.c_int64, C.c_int64]),
    "b2ot_multi_scaling_step": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_void_p,
                                          C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "b2ot_multi_absorb": (C.c_int, [C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_void_p]),
    "b2ot_colsum_static": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p,
                                     C.c_void_p]),
    "b2ot_batch_workspace_bytes": (C.c_size_t, [C.c_int64]),
    "b2ot_batch_config": (C.c_int, [C.c_int, C.c_int, C.c_int64, C.c_int64, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                    C.POINTER(C.c_int)]),
    "b2ot_batch_solve": (C.c_int, [C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_int,
                                   C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double,
                                   C.c_void_p, C.c_size_t, C.POINTER(C.c_int), C.c_int, C.c_void_p]),
    "b2ot_batch_dist": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                  C.c_double, C.c_void_p, C.c_double, C.c_void_p, C.c_void_p]),
    "b2ot_batch_reduce_scratch_bytes": (C.c_size_t, [C.c_int64, C.c_int64]),
    "b2ot_batch_reduce": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_double,
                                    C.c_double, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2ot_strided_gemm": (C.c_int, [C.c_int] + [C.c_int64] * 6 + [C.c_void_p] + [C.c_int64] * 4 + [C.c_void_p] +
                          [C.c_int64] * 4 + [C.c_void_p] + [C.c_int64] * 2 + [C.c_void_p] + [C.c_int64] * 3 +
                          [C.c_double, C.c_double, C.c_void_p]),
    "b2ot_emp_workspace_bytes": (C.c_size_t, [C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "b2ot_emp_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "b2ot_emp_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int64, C.c_int64, C.c_int, C.c_void_p,
                                  C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double,
                                  C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_double, C.c_int, C.c_void_p]),
    "b2ot_emp_set_operand": (C.c_int, [C.c_void_p, C.c_int]),
    "b2ot_emp_set_allreduce": (C.c_int, [C.c_void_p, ALLREDUCE_FN, C.c_void_p]),
    "b2ot_emp_init": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2ot_emp_iterate": (C.c_int, [C.c_void_p, C.c_int]),
    "b2ot_emp_profile": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "b2ot_emp_poll": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "b2ot_emp_solve": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int),
                                 C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "b2ot_emp_export": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_double), C.c_int,
                                  C.POINTER(C.c_int)]),
    "b2ot_emp_destroy": (C.c_int, [C.c_void_p]),
    "b2ot_dist_scratch_bytes": (C.c_size_t, [C.c_int64, C.c_int64, C.c_int]),
    "b2ot_dist": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_int64, C.c_int64,
                            C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "b2ot_emp_plan_block": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                      C.c_int, C.c_int64, C.c_int64, C.c_double, C.c_int, C.c_void_p, C.c_int64,
                                      C.c_void_p, C.c_void_p]),
    "b2ot_emp_loss_scratch_bytes": (C.c_size_t, [C.c_int64, C.c_int64, C.c_int]),
    "b2ot_emp_loss": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int,
                                C.c_int64, C.c_int64, C.c_double, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2ot_comm_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p), C.c_char_p]),
    "b2ot_comm_free": (C.c_int, [C.c_void_p]),
    "b2ot_comm_open": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "b2ot_comm_close": (C.c_int, [C.c_void_p]),
    "b2ot_comm_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.POINTER(C.c_void_p), C.c_size_t]),
    "b2ot_comm_destroy": (C.c_int, [C.c_void_p]),
    "b2ot_comm_slot_bytes": (C.c_size_t, [C.c_void_p]),
    "b2ot_comm_allreduce": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p]),
    "b2ot_dense_set_comm": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b2ot_dense_set_comm_ll": (C.c_int, [C.c_void_p, C.c_int]),
    "b2ot_emp_set_comm": (C.c_int, [C.c_void_p, C.c_void_p]),
}

ST_COMM_TIMEOUT = 5
IPC_HANDLE_BYTES = 64
METRIC_SQEUCLIDEAN, METRIC_EUCLIDEAN = 0, 1
EMP_AUTO, EMP_SIMT, EMP_TC = 0, 1, 2
BMETRIC = {"sqeuclidean": 0, "euclidean": 1, "minkowski": 2, "kl": 3}
BREDUCE_DOT, BREDUCE_ENTROPY = 0, 1
EMP_OPERAND_F16, EMP_OPERAND_TF32 = 0, 1

_lib = None


class B2otError(RuntimeError):
    pass


def load():
    """Load libb2ot.so (once) and attach prototypes.  Raises if the library was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build the CUDA library first "
            "(`python -c 'import __graft_entry__ as g; g.build()'` or `make -C pot_b200/csrc`). "
            "pot_b200 has no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the build is stale
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        msg = load().b2ot_last_error()
        raise B2otError(f"{what} failed (code {rc}): {msg.decode() if msg else '?'}")


def source_hash():
    """sha256 prefix over the kernel sources in the tree, in the Makefile's order (what b2ot_build_info() reports as src=)."""
    import hashlib
    import re

    d = os.path.join(_HERE, "csrc")
    mk = open(os.path.join(d, "Makefile")).read()
    names = re.search(r"^SRCS\s*:=\s*(.*)$", mk, re.M).group(1).split() + re.search(r"^HDRS\s*:=\s*(.*)$", mk, re.M).group(1).split()
    h = hashlib.sha256()
    for n in names:
        with open(os.path.join(d, n), "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def build_matches_sources():
    info = load().b2ot_build_info().decode()
    return ("src=" + source_hash()) in info, info
