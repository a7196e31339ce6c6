"""ctypes binding of include/tahoe_b200.h -- the host-side mirror used by tests and bench.

The product is libtahoe_b200.so (hand-written sm_100a CUDA behind a C ABI).  This module
only marshals pointers.  If the library is missing it raises: there is no Python, numpy
or torch fallback for any compute entry point.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .synth import DENSE_NODE, SPARSE_NODE

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TAHOE_B200_LIB") or os.path.join(HERE, "lib", "libtahoe_b200.so")

OK, ERR_INVALID, ERR_CUDA, ERR_NOT_APPLICABLE, ERR_IO, ERR_NOMEM = range(6)
AUTO, DIRECT, SHARED_DATA_PER_THREAD, SHARED_DATA, SHARED_FOREST, SPLIT_SHARED_FOREST = -1, 0, 1, 2, 3, 4
STRATEGY_NAMES = {DIRECT: "direct", SHARED_DATA_PER_THREAD: "shared_data_per_thread",
                  SHARED_DATA: "shared_data", SHARED_FOREST: "shared_forest",
                  SPLIT_SHARED_FOREST: "split_shared_forest"}
ALL_STRATEGIES = list(STRATEGY_NAMES)

(EXPORT_HEAP_NODES, EXPORT_INTERLEAVED, EXPORT_TREE_ORDER, EXPORT_ORIG_IDS, EXPORT_QUANT_WORDS,
 EXPORT_QUANT_TABLE, EXPORT_QUANT_OFFSETS, EXPORT_LEAF_ORIGIN, EXPORT_SIMHASH, EXPORT_QUANT_LUT,
 EXPORT_QUANT_LUT_MIN, EXPORT_QUANT_LUT_SCALE, EXPORT_SD_TOP_NODES, EXPORT_SD_BLOCKS,
 EXPORT_SD_ORIG_BOTTOM, EXPORT_SPARSE_BFS, EXPORT_SPARSE_BFS_ORIG, EXPORT_SPARSE_TOP,
 EXPORT_SPARSE_DEEP_BLOCKS, EXPORT_SPARSE_BLOCK_ORIG, EXPORT_SPARSE_BLOCK_OFFSETS) = range(21)


class TahoeError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"tahoe status {status}: {msg}")
        self.status = status


class NotApplicable(TahoeError):
    pass


class ForestParams(C.Structure):
    _fields_ = [("num_nodes", C.c_int), ("depth", C.c_int), ("num_trees", C.c_int),
                ("num_cols", C.c_int), ("algo", C.c_int), ("output", C.c_int),
                ("threshold", C.c_float), ("global_bias", C.c_float), ("strategy", C.c_int),
                ("missing", C.c_float)]


class ForestInfo(C.Structure):
    _fields_ = [("num_trees", C.c_int), ("depth", C.c_int), ("num_cols", C.c_int),
                ("is_sparse", C.c_int), ("num_nodes", C.c_int64), ("node_width_bytes", C.c_int),
                ("quantized", C.c_int), ("rank_bits", C.c_int), ("col_bits", C.c_int),
                ("trees_per_chunk", C.c_int), ("num_chunks", C.c_int), ("bytes_heap", C.c_int64),
                ("bytes_interleaved", C.c_int64), ("bytes_quantized", C.c_int64),
                ("exchanged_nodes", C.c_int64), ("mean_depth", C.c_double), ("device", C.c_int),
                ("sm_count", C.c_int), ("smem_optin_bytes", C.c_int), ("col_shift", C.c_int),
                ("smem_warps", C.c_int), ("smem_staged", C.c_int), ("smem_group", C.c_int),
                ("sd_teams", C.c_int), ("sd_team_warps", C.c_int), ("sd_top_levels", C.c_int),
                ("sd_blocked", C.c_int), ("sparse_cut", C.c_int), ("sparse_deep_walks", C.c_double),
                ("sparse_fid_bits", C.c_int), ("sparse_top_depth", C.c_int), ("sparse_top_active", C.c_int),
                ("sparse_top_trees_per_chunk", C.c_int), ("sparse_top_chunks", C.c_int),
                ("sparse_top_deep_walks", C.c_double), ("smem_cluster", C.c_int)]


class Peaks(C.Structure):
    _fields_ = [("smem_read_gbs", C.c_double), ("smem_random_word_loads_per_s", C.c_double),
                ("l2_read_gbs", C.c_double), ("l2_sector_gathers_per_s", C.c_double),
                ("sm_count", C.c_int), ("l2_bytes", C.c_int), ("persisting_l2_max_bytes", C.c_int)]


_lib = None


def lib() -> C.CDLL:
    """Load the CUDA library; fail loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python -m tahoe_b200.build` "
                          "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, sz = C.c_void_p, C.c_size_t
    L.tahoe_forest_create_dense.argtypes = [vp, C.POINTER(ForestParams), C.POINTER(vp)]
    L.tahoe_forest_create_sparse.argtypes = [vp, vp, vp, C.POINTER(ForestParams), C.POINTER(vp)]
    L.tahoe_forest_create_sparse_classes.argtypes = [vp, vp, vp, C.POINTER(ForestParams), C.c_int, vp, C.POINTER(vp)]
    L.tahoe_predict_classes.argtypes = [vp, C.c_int, vp, sz, vp, vp]
    L.tahoe_forest_destroy.argtypes = [vp]
    L.tahoe_forest_destroy.restype = None
    L.tahoe_forest_info.argtypes = [vp, C.POINTER(ForestInfo)]
    L.tahoe_predict.argtypes = [vp, C.c_int, vp, sz, vp, vp, vp]
    L.tahoe_predict_to.argtypes = [vp, C.c_int, vp, sz, vp, vp, vp, vp]
    L.tahoe_enable_peer_access.argtypes = [C.c_int]
    L.tahoe_gather_buffer_create.argtypes = [sz, C.POINTER(vp), C.c_char_p]
    L.tahoe_gather_buffer_open.argtypes = [C.c_char_p, C.POINTER(vp)]
    L.tahoe_gather_buffer_release.argtypes = [vp, C.c_int]
    L.tahoe_measure_peaks.argtypes = [C.POINTER(Peaks)]
    L.tahoe_predict_host.argtypes = [vp, C.c_int, vp, sz, vp]
    L.tahoe_predict_no_alloc.argtypes = [vp, vp, sz, vp, vp]
    L.tahoe_host_copy_baseline.argtypes = [vp, C.c_int, vp, sz, vp]
    L.tahoe_strategy_applicable.argtypes = [vp, C.c_int, sz]
    L.tahoe_select_strategy.argtypes = [vp, sz, C.POINTER(C.c_double)]
    L.tahoe_last_launch_count.argtypes = [vp]
    L.tahoe_set_profiling.argtypes = [vp, C.c_int]
    L.tahoe_profile_read.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int)]
    L.tahoe_load_model_text.argtypes = [C.c_char_p, C.POINTER(vp), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.tahoe_load_data_text.argtypes = [C.c_char_p, C.POINTER(vp), C.POINTER(C.c_int),
                                       C.POINTER(C.c_int), C.POINTER(C.c_float)]
    L.tahoe_save_model_binary.argtypes = [C.c_char_p, vp, C.c_int, C.c_int]
    L.tahoe_load_model_binary.argtypes = L.tahoe_load_model_text.argtypes
    L.tahoe_save_data_binary.argtypes = [C.c_char_p, vp, C.c_int, C.c_int, C.c_float]
    L.tahoe_load_data_binary.argtypes = L.tahoe_load_data_text.argtypes
    L.tahoe_free_host.argtypes = [vp]
    L.tahoe_free_host.restype = None
    L.tahoe_gen_data_device.argtypes = [C.c_uint64, sz, sz, C.c_int, C.c_float, C.c_float, vp, vp]
    L.tahoe_forest_export.argtypes = [vp, C.c_int, vp, C.POINTER(sz)]
    L.tahoe_last_error.restype = C.c_char_p
    L.tahoe_version.restype = C.c_char_p
    _lib = L
    return L


def _check(rc: int):
    if rc == OK:
        return
    msg = lib().tahoe_last_error().decode(errors="replace")
    raise (NotApplicable if rc == ERR_NOT_APPLICABLE else TahoeError)(rc, msg)


def _devptr(t):
    """torch CUDA tensor / int / None -> void*"""
    if t is None:
        return None
    if isinstance(t, int):
        return C.c_void_p(t)
    return C.c_void_p(t.data_ptr())


def _stream_ptr(stream):
    if stream is None:
        try:
            import torch
            if torch.cuda.is_available():
                return C.c_void_p(torch.cuda.current_stream().cuda_stream)
        except ImportError:
            pass
        return None
    return C.c_void_p(stream if isinstance(stream, int) else stream.cuda_stream)


class Forest:
    """Owns a tahoe_forest_t handle (mirror of dense_adaptive_forest, Struct.h:1710)."""

    def __init__(self, handle, num_trees, num_cols):
        self._h = handle
        self.num_trees = num_trees
        self.num_cols = num_cols

    @classmethod
    def from_dense(cls, nodes: np.ndarray, num_trees: int, depth: int, num_cols: int, *,
                   missing: float = 0.0, output: int = 0, threshold: float = 0.0,
                   global_bias: float = 0.0) -> "Forest":
        nodes = np.ascontiguousarray(nodes, dtype=DENSE_NODE)
        assert nodes.shape[0] == num_trees * ((1 << (depth + 1)) - 1), "node count != T*(2^(d+1)-1)"
        ps = ForestParams(0, depth, num_trees, num_cols, 0, output, threshold, global_bias, 0, missing)
        h = C.c_void_p()
        _check(lib().tahoe_forest_create_dense(nodes.ctypes.data_as(C.c_void_p), C.byref(ps), C.byref(h)))
        return cls(h, num_trees, num_cols)

    @classmethod
    def from_sparse(cls, nodes: np.ndarray, tree_offsets: np.ndarray, num_cols: int, *,
                    weights=None, missing: float = 0.0, output: int = 0, threshold: float = 0.0,
                    global_bias: float = 0.0) -> "Forest":
        nodes = np.ascontiguousarray(nodes, dtype=SPARSE_NODE)
        offs = np.ascontiguousarray(tree_offsets, dtype=np.int64)
        T = offs.shape[0] - 1
        w = None if weights is None else np.ascontiguousarray(weights, dtype=np.float32)
        ps = ForestParams(int(offs[-1]), 0, T, num_cols, 0, output, threshold, global_bias, 0, missing)
        h = C.c_void_p()
        _check(lib().tahoe_forest_create_sparse(
            nodes.ctypes.data_as(C.c_void_p), None if w is None else w.ctypes.data_as(C.c_void_p),
            offs.ctypes.data_as(C.c_void_p), C.byref(ps), C.byref(h)))
        return cls(h, T, num_cols)

    @classmethod
    def from_sparse_classes(cls, nodes: np.ndarray, tree_offsets: np.ndarray, num_cols: int, tree_class, num_classes: int,
                            *, weights=None, missing: float = 0.0, output: int = 0, threshold: float = 0.0,
                            global_bias: float = 0.0) -> "Forest":
        """One packed forest for a K-class ensemble (tahoe_forest_create_sparse_classes): tree t scores class tree_class[t]."""
        nodes = np.ascontiguousarray(nodes, dtype=SPARSE_NODE)
        offs = np.ascontiguousarray(tree_offsets, dtype=np.int64)
        tc = np.ascontiguousarray(tree_class, dtype=np.int32)
        T = offs.shape[0] - 1
        assert tc.shape[0] == T
        w = None if weights is None else np.ascontiguousarray(weights, dtype=np.float32)
        ps = ForestParams(int(offs[-1]), 0, T, num_cols, 0, output, threshold, global_bias, 0, missing)
        h = C.c_void_p()
        _check(lib().tahoe_forest_create_sparse_classes(
            nodes.ctypes.data_as(C.c_void_p), None if w is None else w.ctypes.data_as(C.c_void_p),
            offs.ctypes.data_as(C.c_void_p), C.byref(ps), num_classes, tc.ctypes.data_as(C.c_void_p), C.byref(h)))
        f = cls(h, T, num_cols)
        f.num_classes = num_classes
        return f

    def predict_classes(self, data, out=None, *, strategy: int = AUTO, stream=None):
        """-> float32 [num_classes, rows]: one score per class (tahoe_predict_classes; one quantise pass for all classes)."""
        import torch
        rows = self._check_device_args(data, None, None)
        K = self.num_classes
        if out is None:
            out = torch.empty((K, rows), dtype=torch.float32, device=data.device)
        if not (out.is_cuda and out.dtype == torch.float32 and out.is_contiguous() and out.numel() >= K * rows):
            raise ValueError("out must be a contiguous float32 CUDA tensor of num_classes x rows elements")
        _check(lib().tahoe_predict_classes(self._h, strategy, _devptr(data), rows, _devptr(out), _stream_ptr(stream)))
        return out

    def close(self):
        if self._h:
            lib().tahoe_forest_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self) -> ForestInfo:
        inf = ForestInfo()
        _check(lib().tahoe_forest_info(self._h, C.byref(inf)))
        return inf

    def applicable(self, strategy: int, rows: int) -> bool:
        return lib().tahoe_strategy_applicable(self._h, strategy, rows) == OK

    def select(self, rows: int):
        t = (C.c_double * 5)()
        s = lib().tahoe_select_strategy(self._h, rows, t)
        return s, list(t)

    def _check_device_args(self, data, preds, leaf_ids):
        """The C ABI takes bare pointers: a wrong dtype, stride or length is an out-of-bounds access."""
        import torch
        rows = int(data.shape[0])
        if not (data.is_cuda and data.dtype == torch.float32 and data.is_contiguous() and data.dim() == 2
                and data.shape[1] == self.num_cols):
            raise ValueError(f"data must be a contiguous float32 CUDA tensor of shape (rows, {self.num_cols})")
        if preds is not None and not (preds.is_cuda and preds.dtype == torch.float32 and preds.is_contiguous()
                                      and preds.numel() >= rows):
            raise ValueError("preds must be a contiguous float32 CUDA tensor of at least `rows` elements")
        if leaf_ids is not None and not (leaf_ids.is_cuda and leaf_ids.dtype == torch.int32
                                         and leaf_ids.is_contiguous()
                                         and leaf_ids.numel() >= rows * self.num_trees):
            raise ValueError("leaf_ids must be a contiguous int32 CUDA tensor of rows x num_trees elements")
        return rows

    def predict(self, data, preds=None, *, strategy: int = AUTO, leaf_ids=None, stream=None):
        """data / preds / leaf_ids: torch CUDA tensors (float32 [R,C], float32 [R], int32 [R,T])."""
        import torch
        rows = self._check_device_args(data, preds, leaf_ids)
        if preds is None:
            preds = torch.empty(rows, dtype=torch.float32, device=data.device)
        _check(lib().tahoe_predict(self._h, strategy, _devptr(data), rows, _devptr(preds),
                                   _devptr(leaf_ids), _stream_ptr(stream)))
        return preds

    def predict_to(self, data, out, *, partial=None, strategy: int = AUTO, leaf_ids=None, stream=None):
        """tahoe_predict_to: final scores go to `out` (write-only: may be a tensor that lives on ANOTHER
        GPU, mapped through CUDA IPC -- the fused gather of the multi-GPU path); `partial` is local
        scratch for the partial sums between forest chunks (None: owned by the handle)."""
        rows = self._check_device_args(data, partial, leaf_ids)
        import torch
        # (an int is a raw device pointer: a peer-mapped buffer of another GPU, which torch cannot own as a tensor here)
        if not isinstance(out, int) and not (out.is_cuda and out.dtype == torch.float32 and out.is_contiguous()
                                             and out.numel() >= rows):
            raise ValueError("out must be a contiguous float32 CUDA tensor of at least `rows` elements")
        _check(lib().tahoe_predict_to(self._h, strategy, _devptr(data), rows, _devptr(out), _devptr(partial),
                                      _devptr(leaf_ids), _stream_ptr(stream)))
        return out

    def predict_no_alloc(self, data, preds, *, stream=None):
        """tahoe_predict_no_alloc: the shared-memory-forest predict on the library's static device scratch
        (no allocation inside the call; `preds` is the caller's).  Raises TahoeError beyond the row limit."""
        rows = self._check_device_args(data, preds, None)
        _check(lib().tahoe_predict_no_alloc(self._h, _devptr(data), rows, _devptr(preds), _stream_ptr(stream)))
        return preds

    def predict_host(self, data, preds=None, *, strategy: int = AUTO, copy_only: bool = False):
        """Host buffers in, host buffers out (numpy or pinned torch CPU tensors).  copy_only: the same
        slab copies without the kernels (tahoe_host_copy_baseline; preds then holds garbage)."""
        def host_view(x, what, shape_ok):
            if isinstance(x, np.ndarray):
                ok = x.dtype == np.float32 and x.flags["C_CONTIGUOUS"]
                ptr = x.ctypes.data
            else:
                import torch
                ok = (not x.is_cuda) and x.dtype == torch.float32 and x.is_contiguous()
                ptr = x.data_ptr()
            if not (ok and shape_ok(x)):
                raise ValueError(f"{what}: need a C-contiguous float32 host array of the right shape")
            return ptr
        rows = int(data.shape[0])
        dp = host_view(data, "data", lambda x: len(x.shape) == 2 and x.shape[1] == self.num_cols)
        if preds is None:
            preds = np.empty(rows, dtype=np.float32)
        pp = host_view(preds, "preds", lambda x: int(np.prod(x.shape)) >= rows)
        fn = lib().tahoe_host_copy_baseline if copy_only else lib().tahoe_predict_host
        _check(fn(self._h, strategy, C.c_void_p(dp), rows, C.c_void_p(pp)))
        return preds

    def set_profiling(self, on: bool):
        _check(lib().tahoe_set_profiling(self._h, int(on)))

    def profile_read(self):
        """-> (quantize_ms_sum, traverse_ms_sum, predicts_covered)"""
        q, t, n = C.c_double(), C.c_double(), C.c_int()
        _check(lib().tahoe_profile_read(self._h, C.byref(q), C.byref(t), C.byref(n)))
        return q.value, t.value, n.value

    def launches(self) -> int:
        return lib().tahoe_last_launch_count(self._h)

    def export(self, which: int, dtype) -> np.ndarray:
        n = C.c_size_t(0)
        _check(lib().tahoe_forest_export(self._h, which, None, C.byref(n)))
        buf = np.empty(n.value // np.dtype(dtype).itemsize, dtype=dtype)
        _check(lib().tahoe_forest_export(self._h, which, buf.ctypes.data_as(C.c_void_p), C.byref(n)))
        return buf


def measure_peaks() -> dict:
    """L2 / shared-memory read-bandwidth microbenchmarks on the current device (tahoe_measure_peaks)."""
    pk = Peaks()
    _check(lib().tahoe_measure_peaks(C.byref(pk)))
    return {k: getattr(pk, k) for k, _ in Peaks._fields_}


IPC_HANDLE_BYTES = 64


def gather_buffer_create(num_floats: int):
    """-> (device pointer, IPC handle bytes): the owner's gathered-output buffer (tahoe_gather_buffer_create)."""
    p, h = C.c_void_p(), C.create_string_buffer(IPC_HANDLE_BYTES)
    _check(lib().tahoe_gather_buffer_create(num_floats, C.byref(p), h))
    return int(p.value), h.raw


def gather_buffer_open(handle: bytes) -> int:
    """Another process's gathered-output buffer, mapped for kernels of the CURRENT device -> device pointer."""
    p = C.c_void_p()
    _check(lib().tahoe_gather_buffer_open(C.create_string_buffer(handle, IPC_HANDLE_BYTES), C.byref(p)))
    return int(p.value)


def gather_buffer_release(ptr: int, owner: bool) -> None:
    _check(lib().tahoe_gather_buffer_release(C.c_void_p(ptr), int(owner)))


class DevicePointer:
    """A raw device pointer to `size` float32 words as an object torch.as_tensor() understands."""

    def __init__(self, ptr: int, size: int):
        self.__cuda_array_interface__ = {"shape": (size,), "typestr": "<f4", "data": (ptr, False), "version": 2}


def enable_peer_access(peer_device: int) -> None:
    """Kernels on the current device may store into `peer_device`'s memory (tahoe_predict_to's d_out)."""
    _check(lib().tahoe_enable_peer_access(int(peer_device)))


def load_model_text(path: str, binary: bool = False):
    p, t, d = C.c_void_p(), C.c_int(), C.c_int()
    fn = lib().tahoe_load_model_binary if binary else lib().tahoe_load_model_text
    _check(fn(path.encode(), C.byref(p), C.byref(t), C.byref(d)))
    n = t.value * ((1 << (d.value + 1)) - 1)
    arr = np.frombuffer((C.c_char * (n * 12)).from_address(p.value), dtype=DENSE_NODE).copy()
    lib().tahoe_free_host(p)
    return arr, t.value, d.value


def save_model_binary(path: str, nodes: np.ndarray, num_trees: int, depth: int) -> None:
    nodes = np.ascontiguousarray(nodes, dtype=DENSE_NODE)
    assert nodes.size == num_trees * ((1 << (depth + 1)) - 1)
    _check(lib().tahoe_save_model_binary(path.encode(), nodes.ctypes.data_as(C.c_void_p), num_trees, depth))


def load_model_binary(path: str):
    return load_model_text(path, binary=True)


def save_data_binary(path: str, data: np.ndarray, missing: float) -> None:
    data = np.ascontiguousarray(data, dtype=np.float32)
    _check(lib().tahoe_save_data_binary(path.encode(), data.ctypes.data_as(C.c_void_p), data.shape[0],
                                        data.shape[1], C.c_float(missing)))


def load_data_binary(path: str):
    return load_data_text(path, binary=True)


def load_data_text(path: str, binary: bool = False):
    p, r, c, m = C.c_void_p(), C.c_int(), C.c_int(), C.c_float()
    fn = lib().tahoe_load_data_binary if binary else lib().tahoe_load_data_text
    _check(fn(path.encode(), C.byref(p), C.byref(r), C.byref(c), C.byref(m)))
    n = r.value * c.value
    arr = np.frombuffer((C.c_char * (n * 4)).from_address(p.value), dtype=np.float32).copy()
    lib().tahoe_free_host(p)
    return arr.reshape(r.value, c.value), m.value


def gen_data_device(seed: int, rows: int, cols: int, *, row0: int = 0, miss_frac: float = 0.0,
                    missing: float = -999.0, out=None, stream=None):
    import torch
    if out is None:
        out = torch.empty((rows, cols), dtype=torch.float32, device="cuda")
    _check(lib().tahoe_gen_data_device(seed, row0, rows, cols, miss_frac, missing, _devptr(out),
                                       _stream_ptr(stream)))
    return out
