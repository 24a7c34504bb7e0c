"""TEST INFRASTRUCTURE ONLY: ctypes wrapper over oracle/_ref/liboracle.so (the plain-C restatement oracle/oracle.c).

Same method names as oracle.ref.RefSystem so tests can swap one for the other.  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference leg may import this module.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "_ref", "liboracle.so")
_lib = None


def available():
    return os.path.exists(_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_PATH)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_uint, C.c_float, C.c_float, C.c_float, C.c_uint]
        L.orc_last_error.restype = C.c_char_p
        L.orc_last_error.argtypes = [C.c_void_p]
        L.orc_get_timestep.restype = C.c_uint
        L.orc_get_nlist_size.restype = C.c_uint
        vp = C.c_void_p
        L.orc_destroy.argtypes = [vp]
        L.orc_set_threads.argtypes = [vp, C.c_int]
        L.orc_set_particles.argtypes = [vp, vp, vp, vp]
        L.orc_set_accel.argtypes = [vp, vp]
        L.orc_get_particles.argtypes = [vp, vp, vp, vp, vp, vp]
        L.orc_setup_lj.argtypes = [vp, C.c_float, C.c_uint, vp, vp, vp, vp, C.c_int]
        L.orc_set_virial.argtypes = [vp, C.c_int]
        L.orc_setup_integrator.argtypes = [vp, C.c_float]
        L.orc_add_nve.argtypes = [vp, C.c_int, C.c_float, C.c_int]
        L.orc_add_langevin.argtypes = [vp, C.c_float, C.c_uint, vp, C.c_int]
        L.orc_add_nvt.argtypes = [vp, C.c_float, C.c_float]
        L.orc_set_dist_check.argtypes = [vp, C.c_int]
        L.orc_set_kT.argtypes = [vp, C.c_float]
        L.orc_set_box.argtypes = [vp, C.c_float, C.c_float, C.c_float, C.c_int]
        L.orc_get_box.argtypes = [vp, vp]
        L.orc_get_nvt_state.argtypes = [vp, vp]
        L.orc_get_nvt_reservoir_energy.argtypes = [vp]
        L.orc_get_nvt_reservoir_energy.restype = C.c_float
        L.orc_set_timestep.argtypes = [vp, C.c_uint]
        L.orc_get_timestep.argtypes = [vp]
        L.orc_compute_nlist.argtypes = [vp, C.c_uint]
        L.orc_compute_forces.argtypes = [vp, C.c_uint]
        L.orc_run.argtypes = [vp, C.c_uint, C.POINTER(C.c_double)]
        L.orc_step_one.argtypes = [vp, C.c_uint]
        L.orc_step_two.argtypes = [vp, C.c_uint, vp]
        L.orc_get_forces.argtypes = [vp, vp, vp]
        L.orc_get_force_scales.argtypes = [vp, C.c_int, vp]
        L.orc_get_nlist_size.argtypes = [vp]
        L.orc_get_nlist.argtypes = [vp, vp, vp, vp, C.c_uint]
        L.orc_get_stats.argtypes = [vp, vp]
        L.orc_get_cell_dims.argtypes = [vp, vp]
        L.orc_get_cell_list.argtypes = [vp, vp, vp]
        L.orc_get_thermo.argtypes = [vp, vp]
        L.orc_setup_pair.argtypes = [vp, C.c_int, C.c_float, C.c_uint, vp, vp, vp, C.c_int]
        L.orc_set_bonds.argtypes = [vp, C.c_uint, vp, vp, C.c_uint]
        L.orc_setup_bond.argtypes = [vp, C.c_int, vp, vp, vp, vp]
        L.orc_exclusions_from_bonds.argtypes = [vp, C.c_int]
        L.orc_get_force_component.argtypes = [vp, C.c_int, vp, vp]
        L.orc_saru.argtypes = [C.c_uint, C.c_uint, C.c_uint, C.c_uint, vp, vp]
        _lib = L
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleSystem:
    def __init__(self, pos4, box, ntypes=1, vel4=None, image=None, threads=1, n_bond_types=0):
        L = lib()
        self.n_bond_types = int(n_bond_types)
        pos4 = np.ascontiguousarray(pos4, np.float32)
        self.N = pos4.shape[0]
        self.ntypes = ntypes
        self.h = C.c_void_p(L.orc_create(self.N, C.c_float(box[0]), C.c_float(box[1]), C.c_float(box[2]), ntypes))
        L.orc_set_threads(self.h, int(threads))
        self.set_particles(pos4, vel4, image)

    def _check(self, rc):
        if rc:
            raise RuntimeError(lib().orc_last_error(self.h).decode())

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_destroy(self.h)
            self.h = None

    def set_particles(self, pos4=None, vel4=None, image=None):
        a = None if pos4 is None else np.ascontiguousarray(pos4, np.float32)
        b = None if vel4 is None else np.ascontiguousarray(vel4, np.float32)
        c = None if image is None else np.ascontiguousarray(image, np.int32)
        lib().orc_set_particles(self.h, _ptr(a), _ptr(b), _ptr(c))

    def set_accel(self, accel3):
        a = np.ascontiguousarray(accel3, np.float32)
        lib().orc_set_accel(self.h, _ptr(a))

    def setup_lj(self, r_cut, epsilon=1.0, sigma=1.0, alpha=1.0, r_buff=0.4, check_period=1, full=True,
                 shift_mode=0, r_on=None):
        assert full, "the restatement implements the full storage mode (what every GPU class uses)"
        nt = self.ntypes
        bc = lambda x: np.ascontiguousarray(np.broadcast_to(np.asarray(x, np.float64), (nt, nt)))
        eps, sig, alp, rc = bc(epsilon), bc(sigma), bc(alpha), bc(r_cut)
        lj1 = np.ascontiguousarray((4.0 * eps * sig ** 12).astype(np.float32))
        lj2 = np.ascontiguousarray((alp * 4.0 * eps * sig ** 6).astype(np.float32))
        rc = np.ascontiguousarray(rc.astype(np.float32))
        ron = None if r_on is None else np.ascontiguousarray(bc(r_on).astype(np.float32))
        self._check(lib().orc_setup_lj(self.h, C.c_float(r_buff), int(check_period), _ptr(rc), _ptr(lj1), _ptr(lj2),
                                       _ptr(ron), int(shift_mode)))

    def setup_pair(self, kind, r_cut, r_buff=0.4, check_period=1, full=True, shift_mode=0, r_on=None, **coeffs):
        from hoomd_b200 import _lib as B  # coefficient processing only (pure numpy)
        assert full
        nt = self.ntypes
        bc = lambda x: np.ascontiguousarray(np.broadcast_to(np.asarray(x, np.float64), (nt, nt)))
        prm = B.pair_params(kind, nt, **coeffs)
        rc = np.ascontiguousarray(bc(r_cut).astype(np.float32))
        ron = None if r_on is None else np.ascontiguousarray(bc(r_on).astype(np.float32))
        self._check(lib().orc_setup_pair(self.h, B.PAIR_KINDS[kind], C.c_float(r_buff), int(check_period), _ptr(rc),
                                         _ptr(prm), _ptr(ron), int(shift_mode)))

    # ---- BondData + md.bond.* + nlist exclusions (same surface as hoomd_b200.System / oracle.ref.RefSystem) ----
    def set_bonds(self, bonds, types=None, n_bond_types=None):
        b = np.ascontiguousarray(bonds, np.uint32).reshape(-1, 2)
        t = np.zeros(len(b), np.uint32) if types is None else np.ascontiguousarray(types, np.uint32)
        self.n_bond_types = int(n_bond_types or self.n_bond_types or 1)
        self._check(lib().orc_set_bonds(self.h, len(b), _ptr(b), _ptr(t), self.n_bond_types))

    def setup_fene(self, k=30.0, r0=1.5, epsilon=1.0, sigma=1.0):
        nb = self.n_bond_types
        bc = lambda x: np.broadcast_to(np.asarray(x, np.float64), (nb,))
        K = np.ascontiguousarray(bc(k), np.float32)
        R = np.ascontiguousarray(bc(r0), np.float32)
        lj1 = np.ascontiguousarray(4.0 * bc(epsilon) * bc(sigma) ** 12, np.float32)
        lj2 = np.ascontiguousarray(4.0 * bc(epsilon) * bc(sigma) ** 6, np.float32)
        self._check(lib().orc_setup_bond(self.h, 0, _ptr(K), _ptr(R), _ptr(lj1), _ptr(lj2)))

    def setup_harmonic(self, k, r0):
        nb = self.n_bond_types
        K = np.ascontiguousarray(np.broadcast_to(np.asarray(k, np.float64), (nb,)), np.float32)
        R = np.ascontiguousarray(np.broadcast_to(np.asarray(r0, np.float64), (nb,)), np.float32)
        self._check(lib().orc_setup_bond(self.h, 1, _ptr(K), _ptr(R), None, None))

    def exclusions_from_bonds(self, on=True):
        self._check(lib().orc_exclusions_from_bonds(self.h, int(on)))

    def force_component(self, which, virial=False):
        f = np.empty((self.N, 4), np.float32)
        v = np.empty((6, self.N), np.float32) if virial else None
        lib().orc_get_force_component(self.h, int(which), _ptr(f), _ptr(v))
        return f, v

    def set_virial(self, on=True):
        lib().orc_set_virial(self.h, int(on))

    def setup_integrator(self, dt):
        lib().orc_setup_integrator(self.h, C.c_float(dt))

    def add_nve(self, limit=None, zero_force=False):
        lib().orc_add_nve(self.h, int(limit is not None), C.c_float(limit or 0.0), int(zero_force))

    def add_langevin(self, kT, seed, gamma=None, noiseless_t=False):
        g = None if gamma is None else np.ascontiguousarray(np.broadcast_to(gamma, (self.ntypes,)), np.float32)
        lib().orc_add_langevin(self.h, C.c_float(kT), C.c_uint(seed), _ptr(g), int(noiseless_t))

    def set_box(self, Lx, Ly, Lz, scale_particles=True):
        """one BoxResizeUpdater::update towards the orthorhombic box (Lx, Ly, Lz)"""
        lib().orc_set_box(self.h, C.c_float(Lx), C.c_float(Ly), C.c_float(Lz), int(scale_particles))

    def box_lengths(self):
        out = np.zeros(3, np.float32)
        lib().orc_get_box(self.h, _ptr(out))
        return out

    def set_kT(self, kT):
        lib().orc_set_kT(self.h, C.c_float(kT))

    def set_dist_check(self, on=True):
        lib().orc_set_dist_check(self.h, int(on))

    def add_nvt(self, kT, tau):
        lib().orc_add_nvt(self.h, C.c_float(kT), C.c_float(tau))

    def nvt_state(self):
        out = np.zeros(2, np.float32)
        lib().orc_get_nvt_state(self.h, _ptr(out))
        return out

    def nvt_reservoir_energy(self):
        return float(lib().orc_get_nvt_reservoir_energy(self.h))

    def set_timestep(self, ts):
        lib().orc_set_timestep(self.h, int(ts))

    @property
    def timestep(self):
        return lib().orc_get_timestep(self.h)

    def compute_nlist(self, ts=0):
        self._check(lib().orc_compute_nlist(self.h, int(ts)))

    def compute_forces(self, ts=0):
        self._check(lib().orc_compute_forces(self.h, int(ts)))

    def run(self, nsteps):
        el = C.c_double(0)
        self._check(lib().orc_run(self.h, int(nsteps), C.byref(el)))
        return el.value

    def step_one(self, ts):
        lib().orc_step_one(self.h, int(ts))

    def step_two(self, ts, net_force4=None):
        f = None if net_force4 is None else np.ascontiguousarray(net_force4, np.float32)
        lib().orc_step_two(self.h, int(ts), _ptr(f))

    def particles(self):
        N = self.N
        out = dict(pos=np.empty((N, 4), np.float32), vel=np.empty((N, 4), np.float32),
                   accel=np.empty((N, 3), np.float32), image=np.empty((N, 3), np.int32), tag=np.empty(N, np.uint32))
        lib().orc_get_particles(self.h, _ptr(out["pos"]), _ptr(out["vel"]), _ptr(out["accel"]), _ptr(out["image"]),
                                _ptr(out["tag"]))
        return out

    def forces(self, virial=True):
        f = np.empty((self.N, 4), np.float32)
        v = np.empty((6, self.N), np.float32) if virial else None
        lib().orc_get_forces(self.h, _ptr(f), _ptr(v))
        return f, v

    def force_scales(self, which=3):
        """parity-test scales (call after compute_forces): sum_j |f_ij| [N], sum_j |V_ij/2| [N], sum_j |virial term| [6, N]
        over the pair (which & 1) and bond (which & 2) contributions"""
        out = np.zeros((8, self.N), np.float64)
        lib().orc_get_force_scales(self.h, which, _ptr(out))
        return out[0], out[1], out[2:]

    def nlist(self):
        n = np.empty(self.N, np.uint32)
        head = np.empty(self.N, np.uint32)
        size = lib().orc_get_nlist_size(self.h)
        nl = np.empty(size, np.uint32)
        lib().orc_get_nlist(self.h, _ptr(n), _ptr(head), _ptr(nl), size)
        return n, head, nl

    def nlist_stats(self):
        out = np.zeros(4, np.uint64)
        lib().orc_get_stats(self.h, _ptr(out))
        return dict(updates=int(out[0]), forced=int(out[1]), dim=int(out[2]), Nmax0=int(out[3]))

    def cell_list(self):
        d = np.zeros(5, np.uint32)
        lib().orc_get_cell_dims(self.h, _ptr(d))
        dim, nmax = d[:3].copy(), int(d[3])
        ncell = int(dim[0]) * int(dim[1]) * int(dim[2])
        size = np.empty(ncell, np.uint32)
        xyzf = np.empty((ncell, nmax, 4), np.float32)
        lib().orc_get_cell_list(self.h, _ptr(size), _ptr(xyzf))
        return dict(dim=dim, Nmax=nmax, cell_size=size, xyzf=xyzf)

    def thermo(self, ts=None):
        out = np.zeros(5, np.float64)
        lib().orc_get_thermo(self.h, _ptr(out))
        return dict(ke=out[0], pe=out[1], T=out[2], P=out[3], ndof=out[4])


def saru(seed1, seed2, seed3, n=3):
    u = np.empty(n, np.uint32)
    f = np.empty(n, np.float32)
    lib().orc_saru(C.c_uint(seed1), C.c_uint(seed2), C.c_uint(seed3), C.c_uint(n), _ptr(u), _ptr(f))
    return u, f
