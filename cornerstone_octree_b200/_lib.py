"""ctypes binding of libcstone_b200.so (include/cstone_b200.h).

The library is the product; this module only loads it and declares the prototypes.  There is NO fallback:
if the shared object is missing the import fails, and every call checks the returned status and raises."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcstone_b200.so")

MORTON, HILBERT = 0, 1
OPEN, PERIODIC, FIXED = 0, 1, 2


class BoxF32(C.Structure):
    _fields_ = [("limits", C.c_float * 6), ("lengths", C.c_float * 3), ("inv_lengths", C.c_float * 3),
                ("boundaries", C.c_ubyte * 3)]


class BoxF64(C.Structure):
    _fields_ = [("limits", C.c_double * 6), ("lengths", C.c_double * 3), ("inv_lengths", C.c_double * 3),
                ("boundaries", C.c_ubyte * 3)]


class CstoneError(RuntimeError):
    def __init__(self, status, where):
        self.status = status
        super().__init__("%s failed with status %d: %s" % (where, status, status_string(status)))


if not os.path.exists(LIB_PATH):
    raise ImportError(
        "cornerstone_octree_b200: %s not found. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(make -C cornerstone_octree_b200/csrc). There is no CPU fallback." % LIB_PATH)

# libcstone_b200.so links libnccl.so.2.  torch bundles a newer NCCL under the same soname: load torch's copy first
# so that both share one NCCL (loading the system copy first would break `import torch` afterwards).
try:
    import torch  # noqa: F401
except ImportError:  # pure C users of the library link the system NCCL
    pass

lib = C.CDLL(LIB_PATH)

_vp, _sz, _i, _u32, _u64, _i32, _f = C.c_void_p, C.c_size_t, C.c_int, C.c_uint32, C.c_uint64, C.c_int32, C.c_float

lib.cstone_b200_abi_version.restype = _i
lib.cstone_b200_status_string.restype = C.c_char_p
lib.cstone_b200_status_string.argtypes = [_i]
lib.cstone_b200_launch_count.restype = _u64
lib.cstone_b200_host_hilbert_table.restype = C.POINTER(C.c_uint16)
lib.cstone_b200_host_hilbert_table.argtypes = [_i]

_PROTOS = {
    # name: (restype, argtypes)
    "minmax_temp_bytes": (_u64, []),
    "scan_temp_bytes": (_u64, [_u64]),
    "find_neighbors_temp_bytes": (_u64, [_i32]),
    "sequence_u32": (_i, [_vp, _sz, _u32, _vp]),
    "gather": (_i, [_vp, _sz, _vp, _vp, _i, _vp]),
    "scatter": (_i, [_vp, _sz, _vp, _vp, _i, _vp]),
    "gather_multi": (_i, [_vp, _sz, _i, _vp, _vp, _vp, _vp]),
    "exclusive_scan_u32": (_i, [_vp, _vp, _sz, _u32, _vp, _vp]),
    "inclusive_scan_u32": (_i, [_vp, _vp, _sz, _vp, _vp]),
    "exclusive_scan_u32_noalloc": (_i, [_vp, _vp, _sz, _u32, _vp]),
    "upsweep_sum": (_i, [_i32, _vp, _vp, _vp, _vp]),
    "compute_node_layout": (_i, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp]),
}
_PROTOS.update({
    "reduce_temp_bytes": (_u64, []),
    "fill": (_i, [_vp, _sz, _vp, _i, _vp]),
    "scale_f32": (_i, [_vp, _sz, _f, _vp]),
    "scale_f64": (_i, [_vp, _sz, C.c_double, _vp]),
    "increment_u32": (_i, [_vp, _vp, _sz, _u32, _vp]),
    "increment_u64": (_i, [_vp, _vp, _sz, _u64, _vp]),
    "reduce_sum_u32": (_i, [_vp, _sz, _u64, _vp, _vp, _vp]),
    "count_u32": (_i, [_vp, _sz, _u32, _vp, _vp, _vp]),
    "count_u64": (_i, [_vp, _sz, _u64, _vp, _vp, _vp]),
    "lower_bound_u32": (_i, [_vp, _sz, _u32, _vp, _vp, _vp]),
    "lower_bound_u64": (_i, [_vp, _sz, _u64, _vp, _vp, _vp]),
    "lower_bound_i32": (_i, [_vp, _sz, _i32, _vp, _vp, _vp]),
    "lower_bound_i64": (_i, [_vp, _sz, C.c_int64, _vp, _vp, _vp]),
    "lower_bound_f32": (_i, [_vp, _sz, _f, _vp, _vp, _vp]),
    "gather_ranges": (_i, [_vp, _vp, _i, _i32, _vp, _vp, _sz, _i, _vp]),
})
for ks, kt in (("k32", _u32), ("k64", _u64)):
    _PROTOS["rebalance_decision_essential_" + ks] = (_i, [_vp, _vp, _vp, _vp, _vp, kt, kt, _u32, _vp, _i32, _vp])
    _PROTOS["mac_refine_decision_" + ks] = (_i, [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp])
    _PROTOS["protect_ancestors_" + ks] = (_i, [_vp, _vp, _vp, _i32, _vp, _vp, _vp])
    _PROTOS["enforce_keys_" + ks] = (_i, [_vp, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp])
    _PROTOS["range_count_" + ks] = (_i, [_vp, _i32, _vp, _vp, _vp, _i32, _vp, _vp])
    for ts in ("f32", "f64"):
        _PROTOS["geo_mac_spheres_%s_%s" % (ks, ts)] = (_i, [_i, _vp, _i32, _vp, _f, _vp, _vp])
        _PROTOS["set_mac_%s_%s" % (ks, ts)] = (_i, [_i, _vp, _i32, _vp, _f, _vp, _vp])
        _PROTOS["mark_macs_%s_%s" % (ks, ts)] = (_i, [_i, _vp, _vp, _vp, _vp, _vp, _i32, _i, _vp, _vp])
for ts in ("f32", "f64"):
    _PROTOS["move_centers_" + ts] = (_i, [_vp, _i32, _vp, _vp])
    _PROTOS["upsweep_centers_" + ts] = (_i, [_i32, _vp, _vp, _vp, _vp])
for tt in ("f64_f64_f64", "f64_f32_f64", "f32_f32_f32"):
    _PROTOS["compute_leaf_source_center_" + tt] = (_i, [_vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp])
for ks, kt in (("k32", _u32), ("k64", _u64)):
    _PROTOS["focus_tree_create_" + ks] = (_vp, [_i, _u32, _f])
    _PROTOS["focus_tree_destroy_" + ks] = (None, [_vp])
    _PROTOS["focus_tree_view_" + ks] = (_i, [_vp, _vp])
    _PROTOS["focus_tree_set_leaves_" + ks] = (_i, [_vp, _vp, _i32, _vp])
    for ts in ("f32", "f64"):
        _PROTOS["focus_tree_mac_refine_%s_%s" % (ks, ts)] = (_i, [_vp, _vp, kt, kt, kt, kt, _f, _vp, _vp])
    for ts in ("f32", "f64"):
        _PROTOS["focus_tree_update_%s_%s" % (ks, ts)] = (_i, [_vp, _vp, _vp, _u64, kt, kt, _vp, _i32, _vp, _vp])
_PROTOS["compute_fixed_groups"] = (_i, [_u32, _u32, _u32, _vp, _vp])
_PROTOS["group_splits_temp_bytes"] = (_u64, [_u32, _u32])
for ks in ("k32", "k64"):
    for tt in ("f64_f64", "f64_f32", "f32_f32"):
        _PROTOS["compute_group_splits_%s_%s" % (ks, tt)] = (
            _i, [_u32, _u32, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _u32, _f, _vp, _u64, _vp, _vp, _u64, _vp])
for ts in ("f32", "f64"):
    _PROTOS["max_norm_square_" + ts] = (_i, [_vp, _vp, _vp, _sz, _vp, _vp, _vp])
for tt in ("f32_f32", "f64_f32", "f64_f64"):
    _PROTOS["segment_max_" + tt] = (_i, [_vp, _vp, _i, _sz, _sz, _vp, _vp])
for ks in ("k32", "k64"):
    _PROTOS["sort_keys_temp_bytes_" + ks] = (_u64, [_u64])
    _PROTOS["sort_keys_" + ks] = (_i, [_vp, _vp, _sz, _vp, _u64, _vp])
    _PROTOS["count_sfc_gaps_" + ks] = (_i, [_vp, _i32, _vp, _vp])
    _PROTOS["fill_sfc_gaps_" + ks] = (_i, [_vp, _i32, _vp, _vp, _vp])
    for its in ("u32", "u64"):
        _PROTOS["lower_bound_many_%s_%s" % (ks, its)] = (_i, [_vp, _sz, _vp, _sz, _vp, _vp])
for ts in ("f32", "f64"):
    _PROTOS["minmax3_" + ts] = (_i, [_vp, _vp, _vp, _sz, _vp, _vp, _vp])
    _PROTOS["halo_radii_" + ts] = (_i, [_vp, _vp, _sz, _f, _vp, _vp])
    _PROTOS["find_neighbors_" + ts] = (_i, [_vp, _vp, _vp, _vp, _u32, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _f, _u32,
                                            _vp, _vp, _vp])
    _PROTOS["find_neighbors_ws_" + ts] = (_i, [_vp, _vp, _vp, _vp, _u32, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _f, _u32,
                                               _vp, _vp, _i32, _vp, _vp])
for ks in ("k32", "k64"):
    _PROTOS["encode_sfc_" + ks] = (_i, [_i, _vp, _vp, _vp, _vp, _sz, _vp])
    _PROTOS["decode_sfc_" + ks] = (_i, [_i, _vp, _vp, _vp, _vp, _sz, _vp])
    _PROTOS["sort_by_key_temp_bytes_" + ks] = (_u64, [_u64])
    _PROTOS["sort_by_key_" + ks] = (_i, [_vp, _vp, _vp, _vp, _sz, _vp, _u64, _i, _vp])
    _PROTOS["compute_node_counts_" + ks] = (_i, [_vp, _vp, _i32, _vp, _vp, _u32, _i, _vp])
    _PROTOS["compute_node_ops_" + ks] = (_i, [_vp, _i32, _vp, _u32, _vp, _vp, _vp, _vp, _vp])
    _PROTOS["rebalance_decision_" + ks] = (_i, [_vp, _i32, _vp, _u32, _vp, _vp])
    _PROTOS["rebalance_tree_" + ks] = (_i, [_vp, _i32, _i32, _vp, _vp, _vp])
    _PROTOS["update_octree_" + ks] = (_i, [_vp, _vp, _u32, _u32, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _vp])
    _PROTOS["compute_octree_" + ks] = (_i, [_vp, _vp, _u32, _u32, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _vp])
    _PROTOS["build_octree_temp_bytes_" + ks] = (_u64, [_i32])
    _PROTOS["build_octree_" + ks] = (_i, [_vp, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _u64, _vp])
    for ts in ("f32", "f64"):
        _PROTOS["compute_sfc_keys_%s_%s" % (ks, ts)] = (_i, [_i, _vp, _vp, _vp, _vp, _sz, _vp, _vp])
        _PROTOS["compute_geo_centers_%s_%s" % (ks, ts)] = (_i, [_i, _vp, _i32, _vp, _vp, _vp, _vp])
        _PROTOS["find_halos_%s_%s" % (ks, ts)] = (_i, [_i, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _vp, _vp])

for _name, (_res, _args) in _PROTOS.items():
    _fn = getattr(lib, "cstone_b200_" + _name)  # AttributeError here = header/library mismatch
    _fn.restype = _res
    _fn.argtypes = _args

EXPORTED = ["abi_version", "status_string", "launch_count", "device_malloc", "device_free", "device_memcpy",
            "device_memset", "device_synchronize", "set_device"] + sorted(_PROTOS)


def status_string(status):
    return lib.cstone_b200_status_string(int(status)).decode()


def call(name, *args):
    """invoke cstone_b200_<name>; raises CstoneError on a non-zero int status"""
    fn = getattr(lib, "cstone_b200_" + name)
    ret = fn(*args)
    if fn.restype is _i and ret != 0:
        raise CstoneError(ret, "cstone_b200_" + name)
    return ret


def launch_count():
    return int(lib.cstone_b200_launch_count())


# ------------------------------------------------------------------------------------------------ domain engine
# include/cstone_b200_domain.h
_i64 = C.c_int64
lib.cstone_b200_nccl_unique_id.restype = _i
lib.cstone_b200_nccl_unique_id.argtypes = [_vp]
lib.cstone_b200_nccl_comm_init.restype = _i
lib.cstone_b200_nccl_comm_init.argtypes = [_vp, _i, _i, C.POINTER(_vp)]
lib.cstone_b200_nccl_comm_destroy.restype = _i
lib.cstone_b200_nccl_comm_destroy.argtypes = [_vp]
_DOMAIN_PROTOS = {
    "create": (_vp, [_i, _i, C.c_uint, C.c_uint, _f, _vp]),
    "destroy": (None, [_vp]),
    "set_comm": (_i, [_vp, _vp]),
    "set_fields_ready_event": (_i, [_vp, _vp]),
    "set_sfc_kind": (_i, [_vp, _i]),
    "set_end_index": (_i, [_vp, _u64]),
    "set_halo_factor": (_i, [_vp, _f]),
    "cells": (_i, [_vp, _vp]),
    "set_halo_mode": (_i, [_vp, _i]),
    "set_halo_push": (_i, [_vp, _i]),
    "let_view": (_i, [_vp] * 9),
    "sync_begin": (_i64, [_vp, _u64, _vp, _vp, _vp, _vp, _i, _vp, _vp]),
    "sync_grav_begin": (_i64, [_vp, _u64, _vp, _vp, _vp, _vp, _i, _vp, _vp]),
    "let_centers": (_i, [_vp, _vp, _vp, _vp]),
    "focus_peer_exchange": (_i, [_vp, _vp, _i]),
    "focus_global_gather": (_i, [_vp, _vp, _i, _vp]),
    "focus_global_extract": (_i, [_vp, _vp, _i, _vp]),
    "global_octree": (_i, [_vp, _vp, _vp, _vp, _vp]),
    "sync_finish": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "exchange_halos": (_i, [_vp, _i, _vp, _vp]),
    "reapply_sync": (_i, [_vp, _i, _vp, _vp, _vp]),
    "info": (_i, [_vp, _vp]),
    "box": (_i, [_vp, _vp]),
    "stage_ms": (_i, [_vp, _vp]),
    "tree_view": (_i, [_vp, _vp]),
    "global_tree": (_i, [_vp, _vp, _vp, _vp]),
    "find_neighbors": (_i, [_vp, _vp, _vp, _vp, _vp, _u32, _vp, _vp]),
    "set_local_world": (_i, [_vp, _vp]),
}
lib.cstone_b200_local_world_create.restype = _vp
lib.cstone_b200_local_world_create.argtypes = [_i]
lib.cstone_b200_local_world_destroy.restype = None
lib.cstone_b200_local_world_destroy.argtypes = [_vp]
DOMAIN_EXPORTED = ["nccl_unique_id", "nccl_comm_init", "nccl_comm_destroy", "local_world_create", "local_world_destroy",
                   "host_assignment_k32",
                   "host_assignment_k64", "host_limit_box_shrinking_f32", "host_limit_box_shrinking_f64",
                   "host_uniform_bins", "host_limit_boundary_shifts_k32", "host_limit_boundary_shifts_k64",
                   "host_update_treelet_k32", "host_update_treelet_k64"]
for ks in ("k32", "k64"):
    for ts in ("f32", "f64"):
        for _name, (_res, _args) in _DOMAIN_PROTOS.items():
            _fn = getattr(lib, "cstone_b200_domain_%s_%s_%s" % (_name, ks, ts))
            _fn.restype = _res
            _fn.argtypes = _args
            DOMAIN_EXPORTED.append("domain_%s_%s_%s" % (_name, ks, ts))
    _fn = getattr(lib, "cstone_b200_host_assignment_" + ks)
    _fn.restype = _i
    _fn.argtypes = [_i, _vp, _vp, _i32, _vp, _vp, _vp]
    _fn = getattr(lib, "cstone_b200_host_limit_boundary_shifts_" + ks)
    _fn.restype = _i
    _fn.argtypes = [_i, _vp, _vp, _vp, _vp, _i32, _vp]
lib.cstone_b200_host_uniform_bins.restype = _i
lib.cstone_b200_host_uniform_bins.argtypes = [_vp, _i32, _i32, _vp, _vp]
for ts in ("f32", "f64"):
    _fn = getattr(lib, "cstone_b200_host_limit_box_shrinking_" + ts)
    _fn.restype = None
    _fn.argtypes = [_vp, _vp, _vp]
