"""ctypes binding of include/reth_replay.h (libreth_replay.so).

The library is built in-tree by `__graft_entry__.build()` / `reth_b200.build`;
there is NO CPU fallback: a missing library or a missing CUDA device raises.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
LIB_DIR = os.path.join(_HERE, "_lib")
LIB_PATH = os.path.join(LIB_DIR, "libreth_replay.so")
SRC = os.path.join(_HERE, "csrc", "replay.cu")
HEADER = os.path.join(ROOT, "include", "reth_replay.h")

RETH_MAX_COLUMNS = 16
HOST, DEVICE = 0, 1
F32, F64, I32, I64 = 0, 1, 2, 3
MIN_SERVICE, MIN_INPROCESS = 0, 1

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-fmad=false",
              "-std=c++17", "-Xcompiler", "-fPIC,-fvisibility=hidden", "-shared"]


class ReplayError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"reth_replay error {code}: {msg}")
        self.code = code


class Config(C.Structure):
    _fields_ = [("capacity", C.c_int64), ("n_columns", C.c_int32),
                ("row_bytes", C.c_int64 * RETH_MAX_COLUMNS), ("max_batch", C.c_int32),
                ("ring_depth", C.c_int32), ("device", C.c_int32), ("min_mode", C.c_int32),
                ("update_cluster", C.c_int32), ("gather_mode", C.c_int32), ("store_rows", C.c_int32),
                ("frame_bytes", C.c_int32), ("frame_capacity", C.c_int32), ("reserved", C.c_int32 * 3),
                ("out_cast", C.c_int32 * RETH_MAX_COLUMNS)]

COL_FRAMES = 0x100                       # RETH_COL_FRAMES


class Stats(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("appended_rows", "sampled_rows", "updated_rows",
                                         "gathered_rows", "kernel_launches", "tail", "size", "hbm_bytes")]


class PipeInfo(C.Structure):
    _fields_ = [("batch", C.c_int32), ("group", C.c_int32), ("n_groups", C.c_int32), ("upd_entries", C.c_int32),
                ("entry_elems", C.c_int64), ("idx_host", C.c_void_p), ("w_host", C.c_void_p),
                ("upd_host", C.c_void_p)]


class IpcExport(C.Structure):
    _fields_ = [("n_columns", C.c_int32), ("ring_depth", C.c_int32), ("max_batch", C.c_int32), ("device", C.c_int32),
                ("row_bytes", C.c_int64 * RETH_MAX_COLUMNS), ("col_handle", (C.c_ubyte * 64) * RETH_MAX_COLUMNS),
                ("idx_handle", C.c_ubyte * 64), ("w_handle", C.c_ubyte * 64)]


class BoardExport(C.Structure):
    _fields_ = [("device", C.c_int32), ("pid", C.c_int32), ("local_ptr", C.c_uint64), ("handle", C.c_ubyte * 64)]


def _sources():
    d = os.path.join(_HERE, "csrc")
    return [os.path.join(d, f) for f in sorted(os.listdir(d))] + [HEADER]


def build(force=False, verbose=False):
    """nvcc -> reth_b200/_lib/libreth_replay.so (sm_100a, cross-compiles without a GPU)."""
    os.makedirs(LIB_DIR, exist_ok=True)
    if not force and os.path.exists(LIB_PATH):
        if all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in _sources()):
            return LIB_PATH
    nvcc = os.environ.get("NVCC", "nvcc")
    if subprocess.call(["which", nvcc], stdout=subprocess.DEVNULL) != 0:
        nvcc = "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH, SRC]
    subprocess.check_call(cmd)
    return LIB_PATH


_vp, _i64, _i32, _f64, _u64 = C.c_void_p, C.c_int64, C.c_int, C.c_double, C.c_uint64
SIGNATURES = {
    "reth_replay_create": [C.POINTER(Config), C.POINTER(_vp)],
    "reth_replay_destroy": [_vp],
    "reth_replay_abi_version": [],
    "reth_replay_clear": [_vp, _vp],
    "reth_replay_batch_ring_resize": [_vp, _i32, _i32],
    "reth_replay_append": [_vp, _vp, _i32, _vp, _i32, _i32, _i64, _f64, _i32, C.POINTER(_i64), _vp],
    "reth_replay_nstep_open": [_vp, _i32, _i32, _f64, _i32, _i32, _i32, _i64],
    "reth_replay_nstep_reset": [_vp, _i32],
    "reth_replay_append_nstep": [_vp, _i32, _vp, _i32, _vp, _i32, _i32, _i64, _f64, _i32, C.POINTER(_i64),
                                 C.POINTER(_i64), _vp],
    "reth_replay_append_at": [_vp, _vp, _i32, _vp, _i32, _vp, _i32, _i32, _i64, _f64, _i32, _vp],
    "reth_replay_update": [_vp, _vp, _i32, _i32, _vp, _i32, _i32, _i64, _f64, _i32, _vp],
    "reth_replay_update_slot": [_vp, _i32, _vp, _i32, _i32, _f64, _i32, _vp],
    "reth_replay_sample": [_vp, _i32, _vp, _i32, _u64, _f64, _i32, _vp],
    "reth_replay_gather": [_vp, _i32, _vp],
    "reth_replay_sample_multi": [_vp, _i32, _i32, _i32, _vp, _u64, _f64, _vp],
    "reth_replay_gather_multi": [_vp, _i32, _i32, _vp],
    "reth_replay_update_slots": [_vp, _i32, _i32, _vp, _i32, _i32, _f64, _i32, _vp],
    "reth_replay_slot_ptrs": [_vp, _i32, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp),
                              C.POINTER(_vp), C.POINTER(C.c_int32)],
    "reth_replay_slot_read": [_vp, _i32, _vp, _vp, _vp, _vp],
    "reth_replay_group_read": [_vp, _i32, _i32, _vp, _vp, _vp],
    "reth_replay_tree_read": [_vp, _vp, _vp, _vp, _vp],
    "reth_replay_tree_read_noalloc": [_vp, _vp, _vp, _vp, _vp],
    "reth_replay_tree_root": [_vp, C.POINTER(_f64), C.POINTER(_f64), _vp],
    "reth_replay_ring_read": [_vp, _i32, _i64, _i64, _vp, _vp],
    "reth_replay_stats_get": [_vp, C.POINTER(Stats)],
    "reth_replay_root_dev": [_vp, C.POINTER(_vp)],
    "reth_replay_sample_global": [_vp, _i32, _vp, _i32, _i32, _vp, _i32, _u64, _f64, _i32, _vp],
    "reth_replay_slot_meta": [_vp, _i32, C.POINTER(_i64), _vp],
    "reth_replay_uniform_note": [_vp, _vp, _i32, _i32, _i64, _i64, _vp],
    "reth_replay_sample_uniform": [_vp, _i32, _i32, _i32, _vp, _i32, _u64, _vp],
    "reth_replay_fifo_push": [_vp, _vp, _i32, _i32, _i64, _vp, _i32, _i32, _i64, _vp],
    "reth_replay_sample_fifo": [_vp, _i32, _i32, _i32, _vp],
    "reth_replay_sampler_state": [_vp, C.POINTER(_i64)],
    "reth_replay_pipe_sampler": [_vp, _i32],
    "reth_replay_pipe_open": [_vp, _i32, _i32, _u64, C.POINTER(PipeInfo)],
    "reth_replay_pipe_close": [_vp],
    "reth_replay_pipe_stream": [_vp, C.POINTER(_vp)],
    "reth_replay_pipe_issue": [_vp, _f64],
    "reth_replay_pipe_wait": [_vp, C.POINTER(C.c_int32)],
    "reth_replay_pipe_next": [_vp, _f64, C.POINTER(C.c_int32)],
    "reth_replay_gather_from": [_vp, _i32, _vp, _i32, _vp],
    "reth_replay_pipe_reserve": [_vp, _i32],
    "reth_replay_append_frames": [_vp, _vp, _i32, _i64, C.POINTER(_i64), _vp],
    "reth_replay_frame_read": [_vp, _i64, _i64, _vp, _vp],
    "reth_replay_pipe_hold": [_vp, _i32],
    "reth_replay_pipe_advance": [_vp, _f64],
    "reth_replay_pipe_ready": [_vp, C.POINTER(_i32)],
    "reth_replay_pipe_update_begin": [_vp, C.POINTER(C.c_int32)],
    "reth_replay_pipe_update_commit": [_vp, _i32, _i64, _i32, _f64, _i32],
    "reth_replay_update_async": [_vp, _vp, _vp, _i32, _i64, _f64, _i32],
    "reth_replay_pipe_drain": [_vp],
    "reth_replay_ipc_export": [_vp, C.POINTER(IpcExport)],
    "reth_ipc_open": [C.POINTER(IpcExport), C.POINTER(_vp)],
    "reth_ipc_slot_ptrs": [_vp, _i32, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp)],
    "reth_ipc_close": [_vp],
    "reth_replay_gather_to": [_vp, _i32, _vp, _i32, _vp],
    "reth_replay_sample_global_multi": [_vp, _i32, _vp, _i32, _i32, _i32, _i32, _vp, _u64, _f64, _vp],
    "reth_replay_slot_quota": [_vp, _i32, _i32, C.POINTER(_i64), _vp],
    "reth_replay_board_export": [_vp, C.POINTER(BoardExport)],
    "reth_replay_board_connect": [_vp, C.POINTER(BoardExport), _i32, _i32],
    "reth_replay_roots_exchange": [_vp, _i32, _vp],
    "reth_replay_all_roots_dev": [_vp, C.POINTER(_vp)],
    "reth_replay_step_loop": [_vp, _i32, _i32, _vp, _vp, _u64, _f64, _f64, _i32, _vp],
    "reth_replay_step_loop_ingest": [_vp, _i32, _i32, _vp, _vp, _u64, _f64, _f64, _i32, _vp, _vp, _i64, _vp],
    "reth_replay_gather_time": [_vp, C.POINTER(_f64), C.POINTER(_i64)],
    "reth_replay_debug_stamps": [_vp, C.POINTER(_i64)],
}

_lib = None


LZ4F_SIGNATURES = {                      # return a byte count (int64) or a negative error
    "reth_lz4f_content_size": [_vp, _i64],
    "reth_lz4f_decompress": [_vp, _i64, _vp, _i64],
    "reth_lz4f_compress_bound": [_i64],
    "reth_lz4f_compress": [_vp, _i64, _vp, _i64],
    "reth_pack_scatter_rows": [_vp, _i64, _i64, _i32, _vp, _vp, _vp],
}


def lib():
    """Load the C-ABI library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(the replay engine is CUDA-only; there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    for name, args in SIGNATURES.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = C.c_int
    L.reth_replay_last_error.argtypes = [_vp]
    L.reth_replay_last_error.restype = C.c_char_p
    for name, args in LZ4F_SIGNATURES.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = C.c_int64
    _lib = L
    return L


def check(handle, rc):
    if rc != 0:
        msg = lib().reth_replay_last_error(handle)
        raise ReplayError(rc, msg.decode() if msg else "?")
