"""ctypes binding of hoomd_b200/libhoomd_b200.so (the C ABI declared in include/hoomd_b200.h).

There is no fallback: if the library is missing, importing any compute entry point raises.  Build it with
``python -c "import __graft_entry__ as g; g.build()"`` or ``make -C hoomd_b200/csrc``.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HOOMD_B200_LIB", os.path.join(_HERE, "libhoomd_b200.so"))  # override: A/B kernel variants
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "hoomd_b200.h")

c_float_p = C.POINTER(C.c_float)
c_uint_p = C.POINTER(C.c_uint)
c_int_p = C.POINTER(C.c_int)
c_double_p = C.POINTER(C.c_double)
vp = C.c_void_p


class Box(C.Structure):
    """b200_box (BoxDim, hoomd/BoxDim.h:568-576)"""
    _fields_ = [("lo", C.c_float * 3), ("hi", C.c_float * 3), ("L", C.c_float * 3), ("Linv", C.c_float * 3),
                ("xy", C.c_float), ("xz", C.c_float), ("yz", C.c_float), ("periodic", C.c_ubyte * 3),
                ("_pad", C.c_ubyte)]


class Dim3(C.Structure):
    _fields_ = [("x", C.c_uint), ("y", C.c_uint), ("z", C.c_uint)]


class CellListArgs(C.Structure):
    _fields_ = [("d_cell_start", vp), ("d_cell_xyzf", vp), ("d_cell_type", vp), ("d_cell_idx", vp),
                ("d_conditions", vp), ("d_cell_count", vp), ("d_particle_cell", vp), ("d_particle_slot", vp),
                ("d_scratch_idx", vp), ("d_pos", vp), ("d_charge", vp), ("N", C.c_uint), ("n_ghost", C.c_uint),
                ("flag_charge", C.c_int), ("flag_type", C.c_int), ("box", Box), ("dim", Dim3),
                ("ghost_width", C.c_float * 3), ("d_ghost_dir", vp), ("d_ghost_shift", vp), ("d_fused_barrier", vp)]


class NlistArgs(C.Structure):
    _fields_ = [("d_nlist", vp), ("d_n_neigh", vp), ("d_last_pos", vp), ("d_conditions", vp), ("d_Nmax", vp),
                ("d_head_list", vp), ("d_pos", vp), ("N", C.c_uint), ("d_cell_start", vp), ("d_cell_xyzf", vp),
                ("d_cell_type", vp), ("dim", Dim3), ("box", Box), ("ghost_width", C.c_float * 3), ("d_r_cut", vp),
                ("d_r_listsq", vp), ("ntypes", C.c_uint), ("half_storage", C.c_int), ("cand_hint", C.c_uint),
                ("image_box", C.POINTER(Box))]


class PairArgs(C.Structure):
    _fields_ = [("d_force", vp), ("d_virial", vp), ("virial_pitch", C.c_uint), ("N", C.c_uint), ("d_pos", vp),
                ("box", Box), ("d_n_neigh", vp), ("d_nlist", vp), ("d_head_list", vp), ("d_rcutsq", vp),
                ("d_ronsq", vp), ("ntypes", C.c_uint), ("shift_mode", C.c_uint), ("compute_virial", C.c_uint),
                ("threads_per_particle", C.c_uint), ("block_size", C.c_uint), ("wrap_skip_margin", C.c_float)]


class DpdThermoArgs(C.Structure):
    _fields_ = [("d_vel", vp), ("d_tag", vp), ("seed", C.c_uint), ("timestep", C.c_uint), ("T", C.c_float),
                ("deltaT", C.c_float)]


class ForceList(C.Structure):
    _fields_ = [("f", vp * 6), ("v", vp * 6), ("vpitch", C.c_uint * 6)]


class LangevinArgs(C.Structure):
    _fields_ = [("d_gamma", vp), ("n_types", C.c_uint), ("use_lambda", C.c_int), ("lambda_", C.c_float),
                ("T", C.c_float), ("timestep", C.c_uint), ("seed", C.c_uint), ("d_sum_bdenergy", vp),
                ("noiseless_t", C.c_int), ("tally", C.c_int)]


class PDataArrays(C.Structure):
    _fields_ = [("pos", vp), ("vel", vp), ("accel", vp), ("image", vp), ("tag", vp), ("net_force", vp),
                ("charge", vp), ("diameter", vp), ("body", vp)]


class BondArgs(C.Structure):
    """b200_bond_args (bond_args_t, hoomd/md/PotentialBondGPU.cuh:26-98)"""
    _fields_ = [("d_force", vp), ("d_virial", vp), ("virial_pitch", C.c_uint), ("N", C.c_uint), ("d_pos", vp),
                ("d_diameter", vp), ("box", Box), ("d_bond_list", vp), ("pitch", C.c_uint), ("d_n_bonds", vp),
                ("n_bond_types", C.c_uint), ("compute_virial", C.c_uint), ("d_tag", vp), ("d_rtag", vp)]


class MgParams(C.Structure):
    """b200_mg_params"""
    _fields_ = [("world", C.c_uint), ("rank", C.c_uint), ("grid", C.c_uint * 3), ("global_box", Box),
                ("N_global", C.c_uint), ("N_local", C.c_uint), ("Ncap", C.c_uint), ("r_ghost", C.c_float),
                ("capacity_factor", C.c_float), ("ntypes", C.c_uint), ("device", C.c_int)]


# every symbol include/hoomd_b200.h declares; tests/test_abi.py checks this list against the header and the .so
SYMBOLS = [
    "b200_box_init", "b200_box_init_lohi", "b200_compute_cell_list", "b200_cell_list_dimensions",
    "b200_cell_list_export_padded", "b200_cell_list_build_adjacency", "b200_nlist_build_head_list",
    "b200_compute_nlist_binned", "b200_nlist_needs_update_check", "b200_compute_lj_forces",
    "b200_compute_pair_forces", "b200_sys_setup_pair",
    "b200_nlist_filter", "b200_compute_fene_bond_forces", "b200_compute_harmonic_bond_forces",
    "b200_sys_set_bonds", "b200_sys_setup_bond_force", "b200_sys_nlist_exclusions_from_bonds",
    "b200_sys_get_force_component",
    "b200_integrator_sum_net_force", "b200_nve_step_one", "b200_nve_step_two", "b200_langevin_hash_seed",
    "b200_langevin_step_two", "b200_saru_draw", "b200_sfc_generate_sorted_order", "b200_sfc_apply_sorted_order",
    "b200_compute_thermo", "b200_compute_thermo_static", "b200_sys_create", "b200_sys_destroy", "b200_sys_last_error", "b200_sys_set_particles",
    "b200_sys_set_accel", "b200_sys_get_particles", "b200_sys_setup_lj", "b200_sys_set_virial",
    "b200_sys_setup_integrator", "b200_sys_add_nve", "b200_sys_add_langevin", "b200_sys_set_sorter",
    "b200_sys_set_dist_check", "b200_sys_set_langevin_options", "b200_sys_set_diameters",
    "b200_sys_get_langevin_reservoir_energy", "b200_sys_set_group", "b200_sys_set_temperature", "b200_sys_set_box", "b200_sys_get_box", "b200_box_resize_particles",
    "b200_sys_add_nvt", "b200_sys_get_nvt_state", "b200_sys_set_nvt_state", "b200_nvt_mtk_step_one", "b200_nvt_mtk_step_two", "b200_nvt_mtk_advance",
    "b200_compute_dpd_thermo_forces", "b200_sys_set_dpd_thermostat",
    "b200_sys_add_npt", "b200_sys_get_npt_state", "b200_sys_set_npt_state", "b200_sys_get_pressure_tensor",
    "b200_npt_mtk_step_one", "b200_npt_mtk_wrap", "b200_npt_mtk_step_two", "b200_npt_mtk_rescale", "b200_compute_thermo_tensor",
    "b200_sys_sort_now", "b200_sys_set_timestep", "b200_sys_get_timestep", "b200_sys_set_tuning",
    "b200_sys_set_pair_mode", "b200_sys_get_pair_mode", "b200_sys_get_stepping", "b200_sys_set_multistep",
    "b200_sys_compute_nlist", "b200_sys_compute_forces", "b200_sys_run", "b200_sys_step_one", "b200_sys_step_two",
    "b200_sys_get_forces", "b200_sys_get_nlist_size", "b200_sys_get_nlist", "b200_sys_get_stats",
    "b200_sys_get_thermo", "b200_sys_time_kernel",
    "b200_sys_create_mg", "b200_sys_mg_export", "b200_sys_mg_connect", "b200_sys_get_n", "b200_sys_set_tags",
    "b200_sys_mg_stats", "b200_sys_mg_timing", "b200_sys_mg_get_ghosts",
]

# enum B200_PAIR_* (hoomd/md/AllPairPotentials.h)
# "dpd": the same evaluator slot with gamma in the second parameter; the engine's thermostat switch uses it
PAIR_KINDS = {"lj": 0, "gauss": 1, "yukawa": 2, "mie": 3, "force_shifted_lj": 4, "morse": 5, "dpd_conservative": 6, "dpd": 6}


def pair_params(kind, nt, **c):
    """ntypes x ntypes x 4 float32 parameter table of evaluator `kind` from user coefficients, processed like
    hoomd/md/pair.py process_coeff (lj :602-609, gauss :674-678, yukawa :865-869, morse :1276-1281,
    force_shifted_lj :1764-1772, mie :2083-2093) in double precision, then cast."""
    import numpy as np
    bc = lambda x: np.broadcast_to(np.asarray(x, np.float64), (nt, nt))
    p = np.zeros((nt, nt, 4), np.float64)
    if kind in ("lj", "force_shifted_lj"):
        eps, sig, alp = bc(c.get("epsilon", 1.0)), bc(c.get("sigma", 1.0)), bc(c.get("alpha", 1.0))
        p[..., 0] = 4.0 * eps * sig ** 12
        p[..., 1] = alp * 4.0 * eps * sig ** 6
    elif kind == "gauss":
        p[..., 0], p[..., 1] = bc(c["epsilon"]), bc(c["sigma"])
    elif kind == "yukawa":
        p[..., 0], p[..., 1] = bc(c["epsilon"]), bc(c["kappa"])
    elif kind == "mie":
        eps, sig, n, m = bc(c["epsilon"]), bc(c["sigma"]), bc(c["n"]), bc(c["m"])
        pref = (n / (n - m)) * (n / m) ** (m / (n - m))
        p[..., 0], p[..., 1], p[..., 2], p[..., 3] = eps * sig ** n * pref, eps * sig ** m * pref, n, m
    elif kind == "morse":
        p[..., 0], p[..., 1], p[..., 2] = bc(c["D0"]), bc(c["alpha"]), bc(c["r0"])
    elif kind == "dpd_conservative":  # (A, gamma = 0), hoomd/md/pair.py:1513-1516
        p[..., 0] = bc(c["A"])
    elif kind == "dpd":               # (A, gamma), hoomd/md/pair.py:1384-1388
        p[..., 0], p[..., 1] = bc(c["A"]), bc(c["gamma"])
    else:
        raise ValueError("unknown pair potential " + str(kind))
    return np.ascontiguousarray(p.astype(np.float32))


_lib = None


def lib():
    """Load the C-ABI library (once) and attach prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("hoomd_b200: %s is missing -- build the CUDA extension first (no CPU fallback exists)"
                           % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    L.b200_box_init.argtypes = [C.POINTER(Box), C.c_float, C.c_float, C.c_float]
    L.b200_box_init.restype = None
    L.b200_box_init_lohi.argtypes = [C.POINTER(Box), C.c_float * 3, C.c_float * 3, C.c_ubyte * 3]
    L.b200_box_init_lohi.restype = None
    L.b200_cell_list_dimensions.argtypes = [C.POINTER(Box), C.c_float, C.c_uint, C.c_float, C.POINTER(Dim3),
                                            C.c_float * 3]
    L.b200_cell_list_dimensions.restype = None
    L.b200_compute_cell_list.argtypes = [C.POINTER(CellListArgs), vp]
    L.b200_cell_list_export_padded.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_uint, C.c_uint, vp]
    L.b200_cell_list_build_adjacency.argtypes = [vp, Dim3, C.c_uint, C.c_uint, vp]
    L.b200_nlist_build_head_list.argtypes = [vp, vp, vp, vp, C.c_uint, C.c_uint, vp]
    L.b200_compute_nlist_binned.argtypes = [C.POINTER(NlistArgs), vp]
    L.b200_nlist_needs_update_check.argtypes = [vp, vp, vp, C.c_uint, C.POINTER(Box), vp, C.c_float, C.c_uint,
                                                C.c_float, C.c_float * 3, C.c_uint, vp]
    L.b200_compute_lj_forces.argtypes = [C.POINTER(PairArgs), vp, vp]
    L.b200_compute_pair_forces.argtypes = [C.c_int, C.POINTER(PairArgs), vp, vp]
    L.b200_sys_setup_pair.argtypes = [vp, C.c_int, C.c_float, C.c_uint, vp, vp, vp, C.c_int]
    L.b200_nlist_filter.argtypes = [vp, vp, vp, vp, vp, C.c_uint, C.c_uint, vp, vp, C.c_uint, vp]
    L.b200_compute_fene_bond_forces.argtypes = [C.POINTER(BondArgs), vp, vp, vp]
    L.b200_compute_harmonic_bond_forces.argtypes = [C.POINTER(BondArgs), vp, vp, vp]
    L.b200_sys_set_bonds.argtypes = [vp, C.c_uint, vp, vp, C.c_uint]
    L.b200_sys_setup_bond_force.argtypes = [vp, C.c_int, vp, vp, vp, vp]
    L.b200_sys_nlist_exclusions_from_bonds.argtypes = [vp, C.c_int]
    L.b200_sys_get_force_component.argtypes = [vp, C.c_int, vp, vp]
    L.b200_integrator_sum_net_force.argtypes = [vp, vp, C.c_uint, C.POINTER(ForceList), C.c_uint, C.c_int, C.c_int,
                                                vp]
    L.b200_nve_step_one.argtypes = [vp, vp, vp, vp, vp, C.c_uint, C.POINTER(Box), C.c_float, C.c_int, C.c_float,
                                    C.c_int, vp]
    L.b200_nve_step_two.argtypes = [vp, vp, vp, C.c_uint, vp, C.c_float, C.c_int, C.c_float, C.c_int, vp]
    L.b200_langevin_hash_seed.argtypes = [C.c_uint]
    L.b200_langevin_hash_seed.restype = C.c_uint
    L.b200_langevin_step_two.argtypes = [vp, vp, vp, vp, vp, vp, C.c_uint, vp, C.POINTER(LangevinArgs), C.c_float,
                                         C.c_uint, vp]
    L.b200_saru_draw.argtypes = [vp, vp, vp, C.c_uint, C.c_uint, vp]
    L.b200_sfc_generate_sorted_order.argtypes = [vp, vp, vp, vp, C.c_uint64, C.POINTER(C.c_uint64), vp, C.c_uint,
                                                 C.POINTER(Box), C.c_uint, vp]
    L.b200_sfc_apply_sorted_order.argtypes = [C.POINTER(PDataArrays), C.POINTER(PDataArrays), vp, vp, C.c_uint, vp]
    L.b200_compute_thermo.argtypes = [vp, vp, C.c_uint, vp, vp, vp, C.c_uint, vp, C.c_uint, vp]
    L.b200_compute_thermo_static.argtypes = [vp, C.c_uint, vp, vp, vp, C.c_uint, vp, C.c_uint, vp]
    # engine
    L.b200_sys_create.argtypes = [C.c_uint, C.POINTER(Box), C.c_uint, C.c_int]
    L.b200_sys_create.restype = vp
    L.b200_sys_destroy.argtypes = [vp]
    L.b200_sys_destroy.restype = None
    L.b200_sys_last_error.argtypes = [vp]
    L.b200_sys_last_error.restype = C.c_char_p
    L.b200_sys_set_particles.argtypes = [vp, vp, vp, vp]
    L.b200_sys_set_accel.argtypes = [vp, vp]
    L.b200_sys_get_particles.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    L.b200_sys_setup_lj.argtypes = [vp, C.c_float, C.c_uint, vp, vp, vp, vp, C.c_int]
    L.b200_sys_set_virial.argtypes = [vp, C.c_int]
    L.b200_sys_setup_integrator.argtypes = [vp, C.c_float]
    L.b200_sys_add_nve.argtypes = [vp, C.c_int, C.c_float, C.c_int]
    L.b200_sys_add_langevin.argtypes = [vp, C.c_float, C.c_uint, vp, C.c_int]
    L.b200_sys_add_nvt.argtypes = [vp, C.c_float, C.c_float]
    L.b200_sys_set_dist_check.argtypes = [vp, C.c_int]
    L.b200_sys_set_langevin_options.argtypes = [vp, C.c_int, C.c_float, C.c_int]
    L.b200_sys_set_diameters.argtypes = [vp, vp]
    L.b200_sys_get_langevin_reservoir_energy.argtypes = [vp, c_double_p]
    L.b200_sys_set_group.argtypes = [vp, vp, C.c_uint]
    L.b200_sys_set_temperature.argtypes = [vp, C.c_float]
    L.b200_sys_set_box.argtypes = [vp, C.c_float, C.c_float, C.c_float, C.c_int]
    L.b200_sys_get_box.argtypes = [vp, vp]
    L.b200_sys_get_box.restype = None
    L.b200_box_resize_particles.argtypes = [vp, vp, C.c_uint, C.POINTER(Box), C.POINTER(Box), C.c_int, vp]
    L.b200_sys_get_nvt_state.argtypes = [vp, vp]
    L.b200_sys_set_nvt_state.argtypes = [vp, vp]
    L.b200_nvt_mtk_step_one.argtypes = [vp, vp, vp, vp, vp, C.c_uint, C.POINTER(Box), vp, C.c_float, vp]
    L.b200_nvt_mtk_step_two.argtypes = [vp, vp, vp, C.c_uint, vp, vp, C.c_float, vp]
    L.b200_nvt_mtk_advance.argtypes = [vp, vp, C.c_uint, vp, vp, C.c_uint, C.c_uint, C.c_float, C.c_float, C.c_float, vp]
    L.b200_compute_dpd_thermo_forces.argtypes = [C.POINTER(PairArgs), vp, C.POINTER(DpdThermoArgs), vp]
    L.b200_sys_set_dpd_thermostat.argtypes = [vp, C.c_float, C.c_uint]
    L.b200_sys_add_npt.argtypes = [vp, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_uint, C.c_int, C.c_int,
                                   C.c_float]
    L.b200_sys_get_npt_state.argtypes = [vp, vp]
    L.b200_sys_set_npt_state.argtypes = [vp, vp]
    L.b200_sys_get_pressure_tensor.argtypes = [vp, vp]
    L.b200_npt_mtk_step_one.argtypes = [vp, vp, vp, vp, C.c_uint, C.c_float, vp, vp, vp, C.c_float, C.c_int, vp]
    L.b200_npt_mtk_wrap.argtypes = [C.c_uint, vp, vp, C.POINTER(Box), vp]
    L.b200_npt_mtk_step_two.argtypes = [vp, vp, vp, C.c_uint, vp, vp, C.c_float, C.c_float, vp]
    L.b200_npt_mtk_rescale.argtypes = [C.c_uint, vp, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, vp]
    L.b200_compute_thermo_tensor.argtypes = [vp, vp, C.c_uint, vp, vp, vp, C.c_uint, vp, C.c_uint, vp]
    L.b200_sys_get_stepping.argtypes = [vp, vp]
    L.b200_sys_set_multistep.argtypes = [vp, C.c_int]
    L.b200_sys_set_sorter.argtypes = [vp, C.c_uint]
    L.b200_sys_sort_now.argtypes = [vp]
    L.b200_sys_set_timestep.argtypes = [vp, C.c_uint]
    L.b200_sys_set_timestep.restype = None
    L.b200_sys_get_timestep.argtypes = [vp]
    L.b200_sys_get_timestep.restype = C.c_uint
    L.b200_sys_set_tuning.argtypes = [vp, C.c_uint, C.c_int]
    L.b200_sys_set_pair_mode.argtypes = [vp, C.c_int, C.c_uint, C.c_uint, C.c_uint, C.c_uint]
    L.b200_sys_get_pair_mode.argtypes = [vp, C.POINTER(C.c_uint)]
    L.b200_sys_compute_nlist.argtypes = [vp, C.c_uint]
    L.b200_sys_compute_forces.argtypes = [vp, C.c_uint]
    L.b200_sys_run.argtypes = [vp, C.c_uint, c_double_p, c_double_p]
    L.b200_sys_step_one.argtypes = [vp, C.c_uint]
    L.b200_sys_step_two.argtypes = [vp, C.c_uint, vp]
    L.b200_sys_get_forces.argtypes = [vp, vp, vp]
    L.b200_sys_get_nlist_size.argtypes = [vp]
    L.b200_sys_get_nlist_size.restype = C.c_uint
    L.b200_sys_get_nlist.argtypes = [vp, vp, vp, vp, C.c_uint]
    L.b200_sys_get_stats.argtypes = [vp, C.POINTER(C.c_ulonglong)]
    L.b200_sys_get_thermo.argtypes = [vp, c_double_p]
    L.b200_sys_time_kernel.argtypes = [vp, C.c_int, C.c_uint, c_double_p]
    L.b200_sys_create_mg.argtypes = [C.POINTER(MgParams)]
    L.b200_sys_create_mg.restype = vp
    L.b200_sys_mg_export.argtypes = [vp, vp, C.POINTER(Box), c_uint_p]
    L.b200_sys_mg_connect.argtypes = [vp, vp]
    L.b200_sys_get_n.argtypes = [vp]
    L.b200_sys_get_n.restype = C.c_uint
    L.b200_sys_set_tags.argtypes = [vp, vp]
    L.b200_sys_mg_stats.argtypes = [vp, C.POINTER(C.c_ulonglong)]
    L.b200_sys_mg_get_ghosts.argtypes = [vp, vp, vp]
    L.b200_sys_mg_timing.argtypes = [vp, C.POINTER(C.c_ulonglong)]
    _lib = L
    return L


def make_box(Lx, Ly, Lz):
    b = Box()
    lib().b200_box_init(C.byref(b), Lx, Ly, Lz)
    return b


def check(rc, what="b200 call"):
    if rc != 0:
        raise RuntimeError("%s failed: CUDA error %d" % (what, rc))
