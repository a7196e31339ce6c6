"""ctypes loaders for the checker libraries.  TEST INFRASTRUCTURE ONLY.

  Oracle  -> oracle/libtahoe_oracle.so  (our C restatement, tahoe_oracle.c)
  RefShim -> oracle/_ref/libtahoe_ref.so (the reference's own host code, ref_shim.cu)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  tahoe_b200/ never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libtahoe_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libtahoe_ref.so")

DENSE_NODE = np.dtype([("weight", "<f4"), ("val", "<f4"), ("bits", "<i4")])
SPARSE_NODE = np.dtype([("val", "<f4"), ("bits", "<i4"), ("left_idx", "<i4")])


def build(ref: bool = True) -> None:
    """Compile the checker (gcc) and, where /root/reference exists, the shim."""
    subprocess.run(["make", "-s", "-C", HERE, "oracle"], check=True)
    if ref and os.path.exists("/root/reference/BaseTahoeTest.h"):
        subprocess.run(["make", "-s", "-C", HERE, "ref"], check=True)


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        self.lib = C.CDLL(ORACLE_SO)
        L = self.lib
        L.oracle_predict_dense.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int64,
                                           C.c_int, C.c_float, C.c_int, C.c_float, C.c_float,
                                           C.c_void_p, C.c_void_p]
        L.oracle_predict_dense.restype = None
        L.oracle_predict_sparse.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int64,
                                            C.c_int, C.c_float, C.c_int, C.c_float, C.c_float,
                                            C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_predict_sparse.restype = None
        L.oracle_load_model_text.argtypes = [C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int),
                                             C.POINTER(C.c_int)]
        L.oracle_load_data_text.argtypes = [C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int),
                                            C.POINTER(C.c_int), C.POINTER(C.c_float)]
        L.oracle_free.argtypes = [C.c_void_p]
        L.oracle_dense_to_sparse_tree.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.oracle_dense_to_sparse_tree.restype = C.c_int64

    def predict_dense(self, nodes, num_trees, depth, data, missing, output=0, threshold=0.0,
                      global_bias=0.0, want_leaf_ids=False):
        nodes = np.ascontiguousarray(nodes, dtype=DENSE_NODE)
        data = np.ascontiguousarray(data, dtype=np.float32)
        rows, cols = data.shape
        preds = np.empty(rows, dtype=np.float32)
        lids = np.empty((rows, num_trees), dtype=np.int32) if want_leaf_ids else None
        self.lib.oracle_predict_dense(_ptr(nodes), num_trees, depth, _ptr(data), rows, cols,
                                      missing, output, threshold, global_bias, _ptr(preds),
                                      _ptr(lids) if lids is not None else None)
        return (preds, lids) if want_leaf_ids else preds

    def predict_sparse(self, nodes, tree_offsets, data, missing, output=0, threshold=0.0,
                       global_bias=0.0, want_leaf_ids=False):
        nodes = np.ascontiguousarray(nodes, dtype=SPARSE_NODE)
        offs = np.ascontiguousarray(tree_offsets, dtype=np.int64)
        num_trees = offs.shape[0] - 1
        data = np.ascontiguousarray(data, dtype=np.float32)
        rows, cols = data.shape
        preds = np.empty(rows, dtype=np.float32)
        lids = np.empty((rows, num_trees), dtype=np.int32) if want_leaf_ids else None
        visits = C.c_int64(0)
        self.lib.oracle_predict_sparse(_ptr(nodes), _ptr(offs), num_trees, _ptr(data), rows, cols,
                                       missing, output, threshold, global_bias, _ptr(preds),
                                       _ptr(lids) if lids is not None else None,
                                       C.addressof(visits))
        self.last_visits = visits.value
        return (preds, lids) if want_leaf_ids else preds

    def load_model_text(self, path):
        p, t, d = C.c_void_p(), C.c_int(), C.c_int()
        rc = self.lib.oracle_load_model_text(path.encode(), C.byref(p), C.byref(t), C.byref(d))
        if rc != 0:
            raise OSError(f"oracle_load_model_text({path}) -> {rc}")
        n = t.value * ((1 << (d.value + 1)) - 1)
        arr = np.frombuffer((C.c_char * (n * 12)).from_address(p.value), dtype=DENSE_NODE).copy()
        self.lib.oracle_free(p)
        return arr, t.value, d.value

    def load_data_text(self, path):
        p, r, c, m = C.c_void_p(), C.c_int(), C.c_int(), C.c_float()
        rc = self.lib.oracle_load_data_text(path.encode(), C.byref(p), C.byref(r), C.byref(c),
                                            C.byref(m))
        if rc != 0:
            raise OSError(f"oracle_load_data_text({path}) -> {rc}")
        n = r.value * c.value
        arr = np.frombuffer((C.c_char * (n * 4)).from_address(p.value), dtype=np.float32).copy()
        self.lib.oracle_free(p)
        return arr.reshape(r.value, c.value), m.value


class RefShim:
    """The reference's own CPU code.  Exists only where oracle/_ref was built
    (this container, or shipped prebuilt to the GPU box)."""

    @staticmethod
    def available() -> bool:
        return os.path.exists(REF_SO)

    def __init__(self):
        if not os.path.exists(REF_SO):
            build(ref=True)
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO)
        self.lib = C.CDLL(REF_SO)
        L = self.lib
        L.ref_predict.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_long, C.c_int,
                                  C.c_float, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_int]
        L.ref_predict.restype = None
        L.ref_load_model.argtypes = [C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int),
                                     C.POINTER(C.c_int)]
        L.ref_load_data.argtypes = [C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int),
                                    C.POINTER(C.c_int), C.POINTER(C.c_float)]
        L.ref_free.argtypes = [C.c_void_p]
        L.ref_pack.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]

    def predict(self, nodes, num_trees, depth, data, missing, output=0, threshold=0.0,
                global_bias=0.0, nthreads=1):
        nodes = np.ascontiguousarray(nodes, dtype=DENSE_NODE)
        data = np.ascontiguousarray(data, dtype=np.float32)
        rows, cols = data.shape
        preds = np.empty(rows, dtype=np.float32)
        self.lib.ref_predict(_ptr(nodes), num_trees, depth, _ptr(data), rows, cols, missing,
                             output, threshold, global_bias, _ptr(preds), nthreads)
        return preds

    def load_model(self, path):
        p, t, d = C.c_void_p(), C.c_int(), C.c_int()
        rc = self.lib.ref_load_model(path.encode(), C.byref(p), C.byref(t), C.byref(d))
        if rc != 0:
            raise OSError(f"ref_load_model({path}) -> {rc}")
        n = t.value * ((1 << (d.value + 1)) - 1)
        arr = np.frombuffer((C.c_char * (n * 12)).from_address(p.value), dtype=DENSE_NODE).copy()
        self.lib.ref_free(p)
        return arr, t.value, d.value

    def load_data(self, path):
        p, r, c, m = C.c_void_p(), C.c_int(), C.c_int(), C.c_float()
        rc = self.lib.ref_load_data(path.encode(), C.byref(p), C.byref(r), C.byref(c), C.byref(m))
        if rc != 0:
            raise OSError(f"ref_load_data({path}) -> {rc}")
        n = r.value * c.value
        arr = np.frombuffer((C.c_char * (n * 4)).from_address(p.value), dtype=np.float32).copy()
        self.lib.ref_free(p)
        return arr.reshape(r.value, c.value), m.value

    def pack(self, nodes, num_trees, depth, num_cols):
        nodes = np.ascontiguousarray(nodes, dtype=DENSE_NODE)
        n = nodes.shape[0]
        vals = np.empty(n, dtype=np.float32)
        bits = np.empty(n, dtype=np.int32)
        bl = self.lib.ref_pack(_ptr(nodes), num_trees, depth, num_cols, _ptr(vals), _ptr(bits))
        return bl, vals, bits
