"""ctypes binding of oracle/_ref/libsdmm_ref_ckpt.so -- the reference's own JSON serializers of the STree
accelerator (`.asdmm` checkpoints), see oracle/ref_checkpoint.cpp.  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libsdmm_ref_ckpt.so")
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libsdmm_ref_ckpt.so missing: run `make -C oracle ref` where "
                               "/root/reference is mounted")
        L = C.CDLL(LIB_PATH)
        vp, fp = C.c_void_p, C.POINTER(C.c_float)
        L.sdmm_ref_tree_create.restype = vp
        L.sdmm_ref_tree_create.argtypes = [fp, fp, C.c_long]
        L.sdmm_ref_tree_destroy.argtypes = [vp]
        L.sdmm_ref_tree_split_to_depth.argtypes = [vp, C.c_int]
        L.sdmm_ref_tree_size.argtypes = [vp]
        L.sdmm_ref_tree_save_json.argtypes = [vp, C.c_char_p]
        L.sdmm_ref_tree_load_json.argtypes = [vp, C.c_char_p]
        L.sdmm_ref_tree_set_leaf_model.argtypes = [vp, C.c_int, C.c_int, fp]
        L.sdmm_ref_tree_get_leaf_model.argtypes = [vp, C.c_int, C.c_int, fp, fp]
        L.sdmm_ref_tree_stored_nodes.argtypes = [vp]
        L.sdmm_ref_tree_export_stored.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_double)]
        L.sdmm_ref_tree_find_stored.argtypes = [vp, C.c_long, fp, C.POINTER(C.c_int32)]
        _lib = L
    return _lib


def _f(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class RefCheckpointTree:
    """The reference's STree<float, 3, SDMMContext> with its own save / load."""

    def __init__(self, aabb_min=(0, 0, 0), aabb_max=(1, 1, 1), capacity=16):
        lo, hi = np.asarray(aabb_min, np.float32), np.asarray(aabb_max, np.float32)
        self.h = lib().sdmm_ref_tree_create(_f(lo), _f(hi), capacity)

    def __del__(self):
        try:
            lib().sdmm_ref_tree_destroy(self.h)
        except Exception:
            pass

    def split_to_depth(self, depth):
        lib().sdmm_ref_tree_split_to_depth(self.h, depth)

    def size(self):
        return int(lib().sdmm_ref_tree_size(self.h))

    def save(self, path):
        if lib().sdmm_ref_tree_save_json(self.h, os.fsencode(path)) != 0:
            raise RuntimeError("reference save_json failed")

    def load(self, path):
        if lib().sdmm_ref_tree_load_json(self.h, os.fsencode(path)) != 0:
            raise RuntimeError("reference from_json failed")

    def set_leaf_model(self, node, joint):
        """joint [32, K] -> set_mean + prepare_vectorized on the leaf's mixture; returns prepare's bool (or -1)."""
        joint = np.ascontiguousarray(joint, np.float32)
        return int(lib().sdmm_ref_tree_set_leaf_model(self.h, node, joint.shape[1], _f(joint)))

    def leaf_model(self, node):
        """(joint [32, K], frame [18, K]) of a node's mixture; None if the node holds no value / no mixture."""
        K = int(lib().sdmm_ref_tree_get_leaf_model(self.h, node, 0, None, None))
        if K <= 0:
            return None
        joint, frame = np.zeros((32, K), np.float32), np.zeros((18, K), np.float32)
        lib().sdmm_ref_tree_get_leaf_model(self.h, node, K, _f(joint), _f(frame))
        return joint, frame

    def stored_nodes(self, n=None):
        """[n, 13]: is_leaf, axis, child0, child1, has value, aabb.min[3], aabb.max[3], idx, depth."""
        total = int(lib().sdmm_ref_tree_stored_nodes(self.h))
        n = total if n is None else min(n, total)
        out = np.zeros((n, 13), np.float64)
        lib().sdmm_ref_tree_export_stored(self.h, 0, n, out.ctypes.data_as(C.POINTER(C.c_double)))
        return out

    def find(self, pos3):
        pos3 = np.ascontiguousarray(pos3, np.float32)
        out = np.zeros(pos3.shape[1], np.int32)
        lib().sdmm_ref_tree_find_stored(self.h, pos3.shape[1], _f(pos3), out.ctypes.data_as(C.POINTER(C.c_int32)))
        return out
