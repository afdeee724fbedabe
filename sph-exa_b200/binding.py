"""ctypes binding of libsphexa_b200.so (the C ABI declared in include/sphexa_b200.h).

The product path is the CUDA library; there is NO CPU fallback: importing works
without a GPU (so that symbol checks can run), but creating a context raises if
no CUDA device exists, and a missing library raises at import time.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SPHB200_LIB") or os.path.join(_HERE, "libsphexa_b200.so")

F64, MIXED = 0, 1
OPEN, PERIODIC, FIXED = 0, 1, 2

if not os.path.exists(LIB_PATH):
    raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(the sm_100a CUDA library is the only implementation; there is no fallback)")

lib = C.CDLL(LIB_PATH)

vp = C.c_void_p


class Box(C.Structure):
    _fields_ = [("lim", C.c_double * 6), ("len", C.c_double * 3), ("inv", C.c_double * 3), ("bc", C.c_int * 3)]


class TreeView(C.Structure):
    _fields_ = [("numLeafNodes", C.c_int), ("numNodes", C.c_int), ("prefixes", vp), ("childOffsets", vp),
                ("internalToLeaf", vp), ("levelRange", vp), ("leaves", vp), ("layout", vp), ("centers", vp),
                ("sizes", vp)]


FIELD_NAMES = ["x", "y", "z", "h", "m", "vx", "vy", "vz", "x_m1", "y_m1", "z_m1", "temp", "du_m1", "alpha",
               "xm", "kx", "gradh", "prho", "c", "c11", "c12", "c13", "c22", "c23", "c33", "divv", "curlv",
               "ax", "ay", "az", "du", "nc", "dV11", "dV12", "dV13", "dV22", "dV23", "dV33"]
CONSERVED = FIELD_NAMES[:14]
ALWAYS_F64 = {"x", "y", "z", "temp", "du"}


class Fields(C.Structure):
    # n: length of every array (owned + halos); the pair operators pack gather records over [0, n)
    _fields_ = [(n, vp) for n in FIELD_NAMES] + [("n", C.c_size_t)]


class HydroConst(C.Structure):
    _fields_ = [("K", C.c_double), ("ng0", C.c_uint), ("ngmax", C.c_uint), ("Kcour", C.c_double),
                ("Krho", C.c_double), ("gamma", C.c_double), ("muiConst", C.c_double), ("alphamin", C.c_double),
                ("alphamax", C.c_double), ("decay_constant", C.c_double), ("Atmin", C.c_double),
                ("Atmax", C.c_double), ("ramp", C.c_double), ("minDt", C.c_double), ("minDt_m1", C.c_double),
                ("wh", vp), ("whd", vp), ("avClean", C.c_int)]


class TurbConfig(C.Structure):
    _fields_ = [("solWeight", C.c_double), ("Lbox", C.c_double), ("stEnergyPrefac", C.c_double),
                ("stMachVelocity", C.c_double), ("epsilon", C.c_double), ("powerLawExp", C.c_double),
                ("anglesExp", C.c_double), ("stMaxModes", C.c_uint64), ("rngSeed", C.c_uint64), ("stSpectForm", C.c_int)]


class SimConfig(C.Structure):
    _fields_ = [("dtype", C.c_int), ("rank", C.c_int), ("numRanks", C.c_int), ("bucketSize", C.c_uint),
                ("bucketSizeFocus", C.c_uint), ("capacity", C.c_size_t), ("box", Box), ("hc", HydroConst)]


class GroupView(C.Structure):
    """cstone::GroupView (domain/include/cstone/traversal/groups.hpp)"""
    _fields_ = [("firstBody", C.c_uint32), ("lastBody", C.c_uint32), ("numGroups", C.c_uint32),
                ("groupStart", vp), ("groupEnd", vp)]


class OpScope(C.Structure):
    """sphb200_op_scope: the GroupView argument / searchExtFactor of the reference's GPU operators"""
    _fields_ = [("active", vp), ("searchExtFactor", C.c_float), ("dtParticle", vp), ("gatherMask", vp)]


class BdtInfo(C.Structure):
    """sphb200_bdt_info: sph::Timestep (sph/include/sph/timestep.h:42-66) + group counts"""
    _fields_ = [("nextDt", C.c_float), ("elapsedDt", C.c_float), ("totDt", C.c_float), ("numRungs", C.c_int),
                ("substep", C.c_int), ("rungRanges", C.c_uint32 * 5), ("dt_m1", C.c_float * 4),
                ("dt_drift", C.c_float * 4), ("numGroups", C.c_uint32), ("numActiveGroups", C.c_uint32),
                ("searchExtFactor", C.c_float), ("isSynced", C.c_int)]


class RestartBdt(C.Structure):
    """sphb200_restart_bdt: ts::numRungs, ts::dt_m1 and the host pointer of the rung bytes"""
    _fields_ = [("numRungs", C.c_int), ("dt_m1", C.c_float * 4), ("rung", vp)]


class RestartInfo(C.Structure):
    _fields_ = [("iteration", C.c_uint64), ("numParticlesGlobal", C.c_uint64), ("numParticles", C.c_uint64),
                ("dtype", C.c_int), ("hasTurbulence", C.c_int), ("time", C.c_double), ("minDt", C.c_double),
                ("minDt_m1", C.c_double), ("box", Box)]


# every symbol include/sphexa_b200.h declares (checked by tests/test_abi.py)
SYMBOLS = [
    "sphb200_create", "sphb200_destroy", "sphb200_invalidate_records", "sphb200_set_table_variant", "sphb200_last_error", "sphb200_sync", "sphb200_launch_count",
    "sphb200_make_box", "sphb200_sfc_keys", "sphb200_minmax3", "sphb200_sort_by_key", "sphb200_gather",
    "sphb200_node_counts", "sphb200_update_octree", "sphb200_compute_octree", "sphb200_rebalance_decision",
    "sphb200_build_octree", "sphb200_node_centers", "sphb200_layout", "sphb200_neighbors_size", "sphb200_neighbors_from_rowmajor",
    "sphb200_find_neighbors", "sphb200_neighbors_to_rowmajor", "sphb200_xmass", "sphb200_ve_def_gradh",
    "sphb200_eos", "sphb200_iad_divv_curlv", "sphb200_av_switches", "sphb200_momentum_energy", "sphb200_timestep",
    "sphb200_timestep_device_ptr", "sphb200_positions_update_h", "sphb200_pack_ranges", "sphb200_unpack_ranges",
    "sphb200_sim_create", "sphb200_sim_destroy", "sphb200_sim_upload", "sphb200_sim_download", "sphb200_sim_sync",
    "sphb200_sim_compute_forces", "sphb200_sim_integrate", "sphb200_sim_step", "sphb200_sim_forces_phase",
    "sphb200_sim_start_index", "sphb200_sim_end_index", "sphb200_sim_n_with_halos", "sphb200_sim_num_leaves",
    "sphb200_sim_num_global_leaves", "sphb200_sim_min_dt", "sphb200_sim_ttot", "sphb200_sim_total_neighbors",
    "sphb200_pack_ranges_peers", "sphb200_unpack_ranges_peers", "sphb200_sim_step_host", "sphb200_sim_conserved", "sphb200_sim_attach_turbulence", "sphb200_stirring", "sphb200_turb_create",
    "sphb200_turb_destroy", "sphb200_turb_num_modes", "sphb200_turb_get", "sphb200_turb_set_state",
    "sphb200_turb_rng_state", "sphb200_turb_drive", "sphb200_turb_advance", "sphb200_turb_stir", "sphb200_conserved_quantities",
    "sphb200_conserved_device_ptr", "sphb200_sim_fields", "sphb200_sim_tree", "sphb200_sim_keys", "sphb200_sim_neighbors",
    "sphb200_sim_phase_times", "sphb200_sim_set_layout", "sphb200_sim_local_timestep", "sphb200_sim_global_leaves",
    "sphb200_sim_global_counts", "sphb200_sim_leaf_counts", "sphb200_sim_get_box", "sphb200_sim_set_dt",
    "sphb200_segment_max", "sphb200_mark_halo_leaves",
    "sphb200_sim_get_config", "sphb200_sim_set_constants", "sphb200_sim_set_time", "sphb200_sim_set_box",
    "sphb200_sim_turbulence", "sphb200_sim_save", "sphb200_restart_write", "sphb200_restart_open", "sphb200_restart_close",
    "sphb200_restart_get_info", "sphb200_restart_attribute", "sphb200_restart_read_fields", "sphb200_restart_apply",
    "sphb200_compute_groups", "sphb200_group_divv_timestep", "sphb200_group_acc_timestep",
    "sphb200_group_min_timestep", "sphb200_store_rung", "sphb200_extract_groups", "sphb200_targets_from_groups",
    "sphb200_set_scope", "sphb200_mark_groups", "sphb200_sort_group_dt", "sphb200_sequence", "sphb200_rung_ranges",
    "sphb200_drift_positions", "sphb200_compute_positions_groups", "sphb200_update_h_groups",
    "sphb200_sim_bdt_enable", "sphb200_sim_bdt_compute_forces", "sphb200_sim_bdt_integrate", "sphb200_sim_bdt_step",
    "sphb200_sim_bdt_info", "sphb200_sim_bdt_rungs", "sphb200_sim_bdt_group_dt", "sphb200_sim_bdt_active",
    "sphb200_sim_bdt_groups", "sphb200_sim_bdt_restore", "sphb200_sim_bdt_upload_rungs", "sphb200_sim_bdt_download_rungs",
    "sphb200_restart_write_bdt", "sphb200_restart_read_bytes", "sphb200_assemble_cuboid", "sphb200_cut_sphere", "sphb200_hilbert_table", "sphb200_span_sfc_range", "sphb200_momentum_energy_range", "sphb200_av_switches_courant", "sphb200_update_h_into", "sphb200_mach_square_sum", "sphb200_mach_device_ptr",
    "sphb200_sim_mach_rms", "sphb200_layout_noalloc",
    "sphb200_comm_create_loopback", "sphb200_comm_unique_id", "sphb200_comm_create_nccl", "sphb200_comm_destroy",
    "sphb200_comm_num_ranks", "sphb200_comm_is_loopback", "sphb200_domain_create", "sphb200_domain_sync",
    "sphb200_domain_exchange_halos", "sphb200_domain_compute_forces", "sphb200_domain_integrate", "sphb200_domain_step",
    "sphb200_domain_conserved", "sphb200_domain_set_halo_factor", "sphb200_domain_info", "sphb200_sfc_assignment",
    "sphb200_initial_global_tree", "sphb200_find_halos", "sphb200_halo_radii", "sphb200_domain_step_host",
    "sphb200_min_mac_r2", "sphb200_mark_macs", "sphb200_rebalance_decision_essential", "sphb200_enforce_keys",
    "sphb200_protect_ancestors", "sphb200_find_peers_mac", "sphb200_focus_create", "sphb200_focus_destroy",
    "sphb200_focus_update", "sphb200_focus_update_tree", "sphb200_focus_update_min_mac", "sphb200_focus_update_counts",
    "sphb200_focus_set_leaf_counts", "sphb200_focus_get_view", "sphb200_domain_update_focus",
    "sphb200_domain_focus_view", "sphb200_sort_pairs",
]

lib.sphb200_last_error.restype = C.c_char_p
lib.sphb200_launch_count.restype = C.c_uint64
lib.sphb200_launch_count.argtypes = [vp]
lib.sphb200_set_table_variant.argtypes = [vp, C.c_int]
lib.sphb200_neighbors_size.restype = C.c_size_t
lib.sphb200_neighbors_size.argtypes = [C.c_size_t, C.c_uint]
lib.sphb200_timestep_device_ptr.restype = vp
lib.sphb200_conserved_device_ptr.restype = vp
lib.sphb200_conserved_device_ptr.argtypes = [vp]
lib.sphb200_mach_device_ptr.restype = vp
lib.sphb200_mach_device_ptr.argtypes = [vp]
lib.sphb200_sim_mach_rms.argtypes = [vp, C.c_uint64, vp]
lib.sphb200_timestep_device_ptr.argtypes = [vp]
for _n in ("sphb200_sim_start_index", "sphb200_sim_end_index", "sphb200_sim_n_with_halos"):
    getattr(lib, _n).restype = C.c_uint32
    getattr(lib, _n).argtypes = [vp]
for _n in ("sphb200_sim_min_dt", "sphb200_sim_ttot"):
    getattr(lib, _n).restype = C.c_double
    getattr(lib, _n).argtypes = [vp]
for _n in ("sphb200_sim_keys", "sphb200_sim_neighbors", "sphb200_sim_global_leaves", "sphb200_sim_global_counts",
           "sphb200_sim_leaf_counts"):
    getattr(lib, _n).restype = vp
    getattr(lib, _n).argtypes = [vp]
lib.sphb200_sim_total_neighbors.restype = C.c_uint64
lib.sphb200_sim_total_neighbors.argtypes = [vp]
for _n in ("sphb200_sim_num_leaves", "sphb200_sim_num_global_leaves"):
    getattr(lib, _n).argtypes = [vp]
lib.sphb200_destroy.argtypes = [vp]
lib.sphb200_sim_turbulence.restype = vp
lib.sphb200_sim_turbulence.argtypes = [vp]
lib.sphb200_sim_set_time.restype = None
lib.sphb200_sim_set_time.argtypes = [vp, C.c_double]
lib.sphb200_sim_set_box.restype = None
lib.sphb200_sim_set_box.argtypes = [vp, vp]
lib.sphb200_sim_save.argtypes = [vp, C.c_char_p, C.c_uint64, C.c_uint64]
lib.sphb200_restart_write.argtypes = [C.c_char_p, vp, vp, vp, C.c_int, C.c_int, vp]
lib.sphb200_restart_write_bdt.argtypes = [C.c_char_p, vp, vp, vp, C.c_int, C.c_int, vp, vp]
lib.sphb200_restart_read_bytes.argtypes = [vp, C.c_char_p, C.c_uint64, C.c_uint64, vp]
lib.sphb200_restart_open.argtypes = [C.c_char_p, vp]
lib.sphb200_assemble_cuboid.argtypes = [vp, C.c_uint64, C.c_uint64, vp, vp, vp, vp, vp, C.c_size_t, vp, vp, vp, C.c_size_t, vp]
lib.sphb200_cut_sphere.argtypes = [vp, C.c_double, vp, vp, vp, C.c_size_t, vp]
lib.sphb200_restart_close.restype = None
lib.sphb200_restart_close.argtypes = [vp]
lib.sphb200_restart_get_info.argtypes = [vp, vp]
lib.sphb200_restart_attribute.restype = C.c_int64
lib.sphb200_restart_attribute.argtypes = [vp, C.c_char_p, vp, C.c_size_t]
lib.sphb200_restart_read_fields.argtypes = [vp, C.c_uint64, C.c_uint64, C.c_int, vp]
lib.sphb200_restart_apply.argtypes = [vp, vp]
lib.sphb200_compute_groups.argtypes = [vp, C.c_uint32, C.c_uint32, vp, vp, vp, vp, C.c_int, vp, vp, C.c_float, vp,
                                       C.c_uint32, vp]
lib.sphb200_group_divv_timestep.argtypes = [vp, C.c_int, C.c_float, vp, vp, vp]
lib.sphb200_group_acc_timestep.argtypes = [vp, C.c_int, C.c_float, vp, vp, vp, vp, vp]
lib.sphb200_group_min_timestep.argtypes = [vp, C.c_int, vp, vp, vp]
lib.sphb200_store_rung.argtypes = [vp, vp, C.c_uint8, vp]
lib.sphb200_extract_groups.argtypes = [vp, vp, vp, C.c_uint32, C.c_uint32, vp, vp]
lib.sphb200_targets_from_groups.argtypes = [vp, vp, vp, vp]
lib.sphb200_sim_destroy.argtypes = [vp]
lib.sphb200_set_scope.argtypes = [vp, vp]
lib.sphb200_mark_groups.argtypes = [vp, vp, C.c_uint32, vp]
lib.sphb200_sort_group_dt.argtypes = [vp, vp, vp, C.c_uint32]
lib.sphb200_sequence.argtypes = [vp, vp, C.c_uint32, C.c_uint32]
lib.sphb200_rung_ranges.argtypes = [vp, C.c_float, vp, C.c_uint32, C.c_int, C.c_float, vp, vp]
lib.sphb200_drift_positions.argtypes = [vp, C.c_int, vp, C.c_float, C.c_float, vp, vp, vp, vp]
lib.sphb200_compute_positions_groups.argtypes = [vp, C.c_int, vp, C.c_float, vp, vp, vp, vp, vp]
lib.sphb200_update_h_groups.argtypes = [vp, C.c_int, vp, C.c_uint, vp, vp]
lib.sphb200_sim_bdt_enable.argtypes = [vp, C.c_int, C.c_double, C.c_double]
for _n in ("sphb200_sim_bdt_compute_forces", "sphb200_sim_bdt_integrate", "sphb200_sim_bdt_step"):
    getattr(lib, _n).argtypes = [vp]
lib.sphb200_sim_bdt_info.argtypes = [vp, vp]
lib.sphb200_sim_bdt_groups.argtypes = [vp, vp]
lib.sphb200_sim_bdt_restore.argtypes = [vp, C.c_int, vp]
lib.sphb200_sim_bdt_upload_rungs.argtypes = [vp, vp, C.c_size_t]
lib.sphb200_sim_bdt_download_rungs.argtypes = [vp, vp]
for _n in ("sphb200_sim_bdt_rungs", "sphb200_sim_bdt_group_dt", "sphb200_sim_bdt_active"):
    getattr(lib, _n).restype = vp
    getattr(lib, _n).argtypes = [vp]


class SphError(RuntimeError):
    """mirrors the std::runtime_error the reference throws on precondition violations"""


ERR_CAPACITY = 2  # SPHB200_ERR_CAPACITY


def check(rc):
    if rc != 0:
        raise SphError(lib.sphb200_last_error().decode())


def make_box(lim, bc):
    b = Box()
    lib.sphb200_make_box((C.c_double * 6)(*lim), (C.c_int * 3)(*bc), C.byref(b))
    return b


def ptr(t):
    """device pointer of a torch tensor (or None)"""
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous()
    return vp(t.data_ptr())


class Context:
    """owns a sphb200_ctx bound to a CUDA stream (default: torch's current stream)"""

    def __init__(self, device=0, stream=None, private_stream=False):
        import torch
        if not torch.cuda.is_available():
            raise SphError("sph-exa_b200 needs a CUDA device: there is no CPU fallback")
        torch.cuda.set_device(device)
        self.device = device
        if private_stream:
            stream = -1  # SPHB200_PRIVATE_STREAM
        elif stream is None:
            stream = torch.cuda.current_stream(device).cuda_stream
        self.h = vp()
        check(lib.sphb200_create(device, vp(stream), C.byref(self.h)))

    def close(self):
        if self.h:
            lib.sphb200_destroy(self.h)
            self.h = vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        check(lib.sphb200_sync(self.h))

    def setTableVariant(self, variant):
        """kernel-table variant of the pair operators: -1 timed on the data (default), 0 L1, 1 shared memory"""
        check(lib.sphb200_set_table_variant(self.h, C.c_int(variant)))

    @property
    def launches(self):
        return lib.sphb200_launch_count(self.h)
