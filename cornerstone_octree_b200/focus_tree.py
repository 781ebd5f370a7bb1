"""Single-rank focus tree on the device — python mirror of `FocusedOctreeSingleNode<KeyType>`
[ref: focus/octree_focus.hpp:349-427] over the C ABI (`cstone_b200_focus_tree_*`, include/cstone_b200.h)."""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import HILBERT, call


class _View(C.Structure):
    _fields_ = [("num_leaf_nodes", C.c_int32), ("num_internal_nodes", C.c_int32), ("num_nodes", C.c_int32),
                ("reserved", C.c_int32), ("leaves", C.c_void_p), ("leaf_counts", C.c_void_p),
                ("prefixes", C.c_void_p), ("child_offsets", C.c_void_p), ("parents", C.c_void_p),
                ("level_range", C.c_void_p), ("internal_to_leaf", C.c_void_p), ("leaf_to_internal", C.c_void_p),
                ("counts", C.c_void_p), ("macs", C.c_void_p)]


def _copy_out(ptr, n, dtype, device):
    """device pointer owned by the library -> fresh torch tensor"""
    out = torch.empty(n, dtype=dtype, device=device)
    if n:
        torch.cuda.current_stream().synchronize()
        st = _lib.lib.cstone_b200_device_memcpy(C.c_void_p(out.data_ptr()), C.c_void_p(ptr),
                                                C.c_size_t(n * out.element_size()), 2)
        if st != 0:
            raise _lib.CstoneError(st, "cstone_b200_device_memcpy")
    return out


class FocusedOctreeSingleNode:
    def __init__(self, bucket_size, theta, key_dtype=torch.int64, sfc_kind=HILBERT, device="cuda"):
        self.ks = "k32" if key_dtype == torch.int32 else "k64"
        self.kdtype = key_dtype
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("cornerstone_octree_b200: the focus tree lives on a CUDA device")
        self.h = call("focus_tree_create_" + self.ks, sfc_kind, bucket_size, theta)
        if not self.h:
            raise ValueError("focus_tree_create: invalid arguments")

    def close(self):
        if getattr(self, "h", None):
            call("focus_tree_destroy_" + self.ks, self.h)
            self.h = None

    __del__ = close

    def update(self, box, particle_keys, focus_start, focus_end, mandatory_keys=()):
        """one update step; particle_keys: sorted device tensor; mandatory_keys: host sequence.  -> converged"""
        bits = 32 if self.ks == "k32" else 64
        npk = np.uint32 if bits == 32 else np.uint64
        mand = np.ascontiguousarray(np.asarray(list(mandatory_keys), dtype=npk))
        conv = C.c_int32(0)
        ts = "f32" if box.__class__.__name__ == "BoxF32" else "f64"
        mask = (1 << bits) - 1
        call("focus_tree_update_%s_%s" % (self.ks, ts), self.h, C.byref(box), particle_keys.data_ptr(),
             particle_keys.numel(), int(focus_start) & mask, int(focus_end) & mask,
             mand.ctypes.data_as(C.c_void_p) if mand.size else None, mand.size, C.byref(conv),
             torch.cuda.current_stream().cuda_stream)
        return bool(conv.value)

    def set_leaves(self, leaves):
        """replace the leaf array (device tensor, num_leaf_nodes + 1 keys)"""
        call("focus_tree_set_leaves_" + self.ks, self.h, leaves.data_ptr(), leaves.numel() - 1,
             torch.cuda.current_stream().cuda_stream)

    def mac_refine(self, box, old_focus_start, old_focus_end, focus_start, focus_end, inv_theta):
        """one step of macRefineGpu [ref: focus/octree_focus.hpp:318-342] -> converged"""
        mask = (1 << (32 if self.ks == "k32" else 64)) - 1
        conv = C.c_int32(0)
        ts = "f32" if box.__class__.__name__ == "BoxF32" else "f64"
        call("focus_tree_mac_refine_%s_%s" % (self.ks, ts), self.h, C.byref(box), int(old_focus_start) & mask,
             int(old_focus_end) & mask, int(focus_start) & mask, int(focus_end) & mask, inv_theta, C.byref(conv),
             torch.cuda.current_stream().cuda_stream)
        return bool(conv.value)

    def _view(self):
        v = _View()
        call("focus_tree_view_" + self.ks, self.h, C.byref(v))
        return v

    def tree_leaves(self):
        v = self._view()
        return _copy_out(v.leaves, v.num_leaf_nodes + 1, self.kdtype, self.device)

    def leaf_counts(self):
        v = self._view()
        if not v.leaf_counts:
            return torch.empty(0, dtype=torch.int32, device=self.device)
        return _copy_out(v.leaf_counts, v.num_leaf_nodes, torch.int32, self.device)

    def node_arrays(self):
        """(prefixes, child_offsets, counts, macs) of the linked tree"""
        v = self._view()
        return (_copy_out(v.prefixes, v.num_nodes, self.kdtype, self.device),
                _copy_out(v.child_offsets, v.num_nodes, torch.int32, self.device),
                _copy_out(v.counts, v.num_nodes, torch.int32, self.device),
                _copy_out(v.macs, v.num_nodes, torch.int8, self.device))
