"""TEST INFRASTRUCTURE ONLY: ctypes wrapper over oracle/_ref/libhoomd_ref.so.

libhoomd_ref.so is the UNMODIFIED reference (HOOMD-blue 2.3.0 CPU classes, SINGLE_PRECISION) compiled from
/root/reference by oracle/Makefile plus the C harness oracle/ref_harness.cc.  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference leg may import this module.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "_ref", "libhoomd_ref.so")
_lib = None

f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


def available():
    return os.path.exists(_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_PATH, mode=C.RTLD_GLOBAL)
        _lib.ref_create.restype = C.c_void_p
        _lib.ref_create.argtypes = [C.c_uint, C.c_float, C.c_float, C.c_float, C.c_uint]
        _lib.ref_create_bonded.restype = C.c_void_p
        _lib.ref_create_bonded.argtypes = [C.c_uint, C.c_float, C.c_float, C.c_float, C.c_uint, C.c_uint]
        _lib.ref_add_bonds.argtypes = [C.c_void_p, C.c_uint, C.c_void_p, C.c_void_p]
        _lib.ref_setup_bond.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.ref_nlist_exclusions_from_bonds.argtypes = [C.c_void_p, C.c_int]
        _lib.ref_get_bond_forces.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.ref_setup_pair.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_int]
        _lib.ref_add_nvt.argtypes = [C.c_void_p, C.c_float, C.c_float]
        _lib.ref_add_npt.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_uint, C.c_int,
                                     C.c_int, C.c_float]
        _lib.ref_get_npt_state.argtypes = [C.c_void_p, C.c_void_p]
        _lib.ref_get_pressure_tensor.argtypes = [C.c_void_p, C.c_uint, C.c_void_p]
        _lib.ref_set_dpd_thermostat.argtypes = [C.c_void_p, C.c_float, C.c_uint]
        _lib.ref_set_dist_check.argtypes = [C.c_void_p, C.c_int]
        _lib.ref_set_kT.argtypes = [C.c_void_p, C.c_float]
        _lib.ref_set_box.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int]
        _lib.ref_get_box.argtypes = [C.c_void_p, C.c_void_p]
        _lib.ref_get_nvt_state.argtypes = [C.c_void_p, C.c_void_p]
        _lib.ref_get_nvt_reservoir_energy.argtypes = [C.c_void_p, C.c_void_p]
        _lib.ref_last_error.restype = C.c_char_p
        _lib.ref_get_lj_energy.restype = C.c_double
        _lib.ref_get_langevin_reservoir_energy.restype = C.c_double
        _lib.ref_get_langevin_reservoir_energy.argtypes = [C.c_void_p]
        _lib.ref_set_group_tags.argtypes = [C.c_void_p, C.c_void_p, C.c_uint]
        _lib.ref_set_diameters.argtypes = [C.c_void_p, C.c_void_p]
        _lib.ref_cell_compute_tdb.argtypes = [C.c_void_p, C.c_uint]
        _lib.ref_get_cell_tdb.argtypes = [C.c_void_p, C.c_void_p]
        _lib.ref_add_langevin_ex.argtypes = [C.c_void_p, C.c_float, C.c_uint, C.c_void_p, C.c_int, C.c_int, C.c_float,
                                             C.c_int]
        for name in ("ref_destroy", "ref_last_error", "ref_get_N", "ref_get_timestep", "ref_get_virial_pitch",
                     "ref_get_nlist_size", "ref_get_lj_energy", "ref_sort_now"):
            getattr(_lib, name).argtypes = [C.c_void_p]
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class RefSystem:
    """One reference simulation: ParticleData + nlist.cell + pair.lj + integrate.mode_standard + nve/langevin."""

    def __init__(self, pos4, box, ntypes=1, vel4=None, image=None, n_bond_types=0):
        L = lib()
        pos4 = np.ascontiguousarray(pos4, np.float32)
        self.N = pos4.shape[0]
        self.ntypes = ntypes
        self.n_bond_types = int(n_bond_types)
        if n_bond_types:
            self.h = C.c_void_p(L.ref_create_bonded(self.N, C.c_float(box[0]), C.c_float(box[1]), C.c_float(box[2]),
                                                    ntypes, int(n_bond_types)))
        else:
            self.h = C.c_void_p(L.ref_create(self.N, C.c_float(box[0]), C.c_float(box[1]), C.c_float(box[2]), ntypes))
        self._check(0)
        self.set_particles(pos4, vel4, image)

    def _check(self, rc):
        err = lib().ref_last_error(self.h)
        if rc or err:
            raise RuntimeError(err.decode())

    def __del__(self):
        if getattr(self, "h", None):
            lib().ref_destroy(self.h)
            self.h = None

    def set_particles(self, pos4=None, vel4=None, image=None):
        a = None if pos4 is None else np.ascontiguousarray(pos4, np.float32)
        b = None if vel4 is None else np.ascontiguousarray(vel4, np.float32)
        c = None if image is None else np.ascontiguousarray(image, np.int32)
        lib().ref_set_particles(self.h, _ptr(a), _ptr(b), _ptr(c))

    def set_accel(self, accel3):
        a = np.ascontiguousarray(accel3, np.float32)
        lib().ref_set_accel(self.h, _ptr(a))

    def set_periodic(self, px, py, pz):
        lib().ref_set_periodic(self.h, int(px), int(py), int(pz))

    def setup_lj(self, r_cut, epsilon=1.0, sigma=1.0, alpha=1.0, r_buff=0.4, check_period=1, full=True,
                 shift_mode=0, r_on=None):
        """r_cut/epsilon/sigma/alpha: scalars or ntypes x ntypes arrays. lj1/lj2 as hoomd/md/pair.py:602-609."""
        nt = self.ntypes
        bc = lambda x: np.ascontiguousarray(np.broadcast_to(np.asarray(x, np.float64), (nt, nt)))
        eps, sig, alp, rc = bc(epsilon), bc(sigma), bc(alpha), bc(r_cut)
        lj1 = (4.0 * eps * sig ** 12).astype(np.float32)
        lj2 = (alp * 4.0 * eps * sig ** 6).astype(np.float32)
        rc = rc.astype(np.float32)
        ron = None if r_on is None else bc(r_on).astype(np.float32)
        rcode = lib().ref_setup_lj(self.h, C.c_float(r_buff), int(check_period), int(full), _ptr(rc), _ptr(lj1),
                                   _ptr(lj2), _ptr(ron), int(shift_mode))
        self._check(rcode)
        self.lj1, self.lj2, self.rcut = lj1, lj2, rc

    # ---- BondData + md.bond.* + nlist exclusions (same surface as hoomd_b200.System) ----
    def set_bonds(self, bonds, types=None, n_bond_types=None):
        b = np.ascontiguousarray(bonds, np.uint32).reshape(-1, 2)
        t = np.zeros(len(b), np.uint32) if types is None else np.ascontiguousarray(types, np.uint32)
        self._check(lib().ref_add_bonds(self.h, len(b), _ptr(b), _ptr(t)))

    def setup_fene(self, k=30.0, r0=1.5, epsilon=1.0, sigma=1.0):
        nb = self.n_bond_types
        bc = lambda x: np.broadcast_to(np.asarray(x, np.float64), (nb,))
        K = np.ascontiguousarray(bc(k), np.float32)
        R = np.ascontiguousarray(bc(r0), np.float32)
        lj1 = np.ascontiguousarray(4.0 * bc(epsilon) * bc(sigma) ** 12, np.float32)
        lj2 = np.ascontiguousarray(4.0 * bc(epsilon) * bc(sigma) ** 6, np.float32)
        self._check(lib().ref_setup_bond(self.h, 0, _ptr(K), _ptr(R), _ptr(lj1), _ptr(lj2)))

    def setup_harmonic(self, k, r0):
        nb = self.n_bond_types
        K = np.ascontiguousarray(np.broadcast_to(np.asarray(k, np.float64), (nb,)), np.float32)
        R = np.ascontiguousarray(np.broadcast_to(np.asarray(r0, np.float64), (nb,)), np.float32)
        self._check(lib().ref_setup_bond(self.h, 1, _ptr(K), _ptr(R), None, None))

    def exclusions_from_bonds(self, on=True):
        self._check(lib().ref_nlist_exclusions_from_bonds(self.h, int(on)))

    def force_component(self, which, virial=False):
        if which == 0:
            return self.forces(virial)
        f = np.empty((self.N, 4), np.float32)
        v = np.empty((6, self.N), np.float32) if virial else None
        lib().ref_get_bond_forces(self.h, _ptr(f), _ptr(v))
        return f, v

    def setup_pair(self, kind, r_cut, r_buff=0.4, check_period=1, full=True, shift_mode=0, r_on=None, **coeffs):
        """md.pair.<kind> (gauss | yukawa | mie | force_shifted_lj | morse | lj) on the reference's PotentialPair<evaluator>"""
        from hoomd_b200 import _lib as B  # coefficient processing only (pure numpy, hoomd/md/pair.py process_coeff)
        nt = self.ntypes
        bc = lambda x: np.ascontiguousarray(np.broadcast_to(np.asarray(x, np.float64), (nt, nt)))
        prm = B.pair_params(kind, nt, **coeffs)
        rc = np.ascontiguousarray(bc(r_cut).astype(np.float32))
        ron = None if r_on is None else np.ascontiguousarray(bc(r_on).astype(np.float32))
        self._check(lib().ref_setup_pair(self.h, B.PAIR_KINDS[kind], C.c_float(r_buff), int(check_period), int(full),
                                         _ptr(rc), _ptr(prm), _ptr(ron), int(shift_mode)))

    def setup_dpd_thermostat(self, r_cut, A, gamma, kT, seed, r_buff=0.4, check_period=1, full=True):
        """md.pair.dpd(r_cut, nlist, kT, seed) + pair_coeff.set(A=, gamma=): PotentialPairDPDThermoDPD"""
        from hoomd_b200 import _lib as B
        nt = self.ntypes
        bc = lambda x: np.ascontiguousarray(np.broadcast_to(np.asarray(x, np.float64), (nt, nt)))
        prm = B.pair_params("dpd", nt, A=A, gamma=gamma)
        rc = np.ascontiguousarray(bc(r_cut).astype(np.float32))
        self._check(lib().ref_setup_pair(self.h, 7, C.c_float(r_buff), int(check_period), int(full), _ptr(rc), _ptr(prm), None, 0))
        self._check(lib().ref_set_dpd_thermostat(self.h, C.c_float(kT), C.c_uint(seed)))

    def set_virial(self, on=True):
        lib().ref_set_virial(self.h, int(on))

    def setup_integrator(self, dt):
        self._check(lib().ref_setup_integrator(self.h, C.c_float(dt)))

    def add_nve(self, limit=None, zero_force=False):
        self._check(lib().ref_add_nve(self.h, int(limit is not None), C.c_float(limit or 0.0), int(zero_force)))

    def set_group(self, tags=None):
        """the group of the integration methods added AFTER this call: member tags (hoomd.group.tag_list), None = all"""
        t = np.zeros(0, np.uint32) if tags is None else np.ascontiguousarray(tags, np.uint32)
        self._check(lib().ref_set_group_tags(self.h, _ptr(t), C.c_uint(len(t))))

    def set_diameters(self, diameter):
        d = np.ascontiguousarray(diameter, np.float32)
        assert len(d) == self.N
        lib().ref_set_diameters(self.h, _ptr(d))

    def add_langevin_ex(self, kT, seed, gamma=None, noiseless_t=False, dscale=None, tally=False):
        """md.integrate.langevin(..., dscale=lambda, tally=...): gamma = lambda * diameter when dscale is given"""
        g = None if gamma is None else np.ascontiguousarray(gamma, np.float32)
        self._check(lib().ref_add_langevin_ex(self.h, C.c_float(kT), C.c_uint(seed), _ptr(g), C.c_int(bool(noiseless_t)),
                                              C.c_int(dscale is not None), C.c_float(dscale or 0.0), C.c_int(bool(tally))))

    def langevin_reservoir_energy(self):
        return float(lib().ref_get_langevin_reservoir_energy(self.h))

    def add_langevin(self, kT, seed, gamma=None, noiseless_t=False):
        g = None if gamma is None else np.ascontiguousarray(np.broadcast_to(gamma, (self.ntypes,)), np.float32)
        self._check(lib().ref_add_langevin(self.h, C.c_float(kT), C.c_uint(seed), _ptr(g), int(noiseless_t)))

    def set_box(self, Lx, Ly, Lz, scale_particles=True):
        """one BoxResizeUpdater::update towards the orthorhombic box (Lx, Ly, Lz)"""
        self._check(lib().ref_set_box(self.h, C.c_float(Lx), C.c_float(Ly), C.c_float(Lz), int(scale_particles)))

    def box_lengths(self):
        out = np.zeros(3, np.float32)
        lib().ref_get_box(self.h, _ptr(out))
        return out

    def set_kT(self, kT):
        self._check(lib().ref_set_kT(self.h, C.c_float(kT)))

    def set_dist_check(self, on=True):
        self._check(lib().ref_set_dist_check(self.h, int(on)))

    def add_nvt(self, kT, tau):
        self._check(lib().ref_add_nvt(self.h, C.c_float(kT), C.c_float(tau)))

    COUPLE = {"none": 0, "xy": 1, "xz": 2, "yz": 3, "xyz": 4}

    def add_npt(self, kT, tau, P, tauP, couple="xyz", x=True, y=True, z=True, nph=False, rescale_all=False, gamma=0.0):
        """md.integrate.npt on the method group (TwoStepNPTMTK, orthorhombic degrees of freedom)"""
        flags = (1 if x else 0) | (2 if y else 0) | (4 if z else 0)
        self._check(lib().ref_add_npt(self.h, C.c_float(kT), C.c_float(tau), C.c_float(P), C.c_float(tauP),
                                      self.COUPLE[couple], flags, int(nph), int(rescale_all), C.c_float(gamma)))

    def npt_state(self):
        """{eta, xi, nuxx, nuxy, nuxz, nuyy, nuyz, nuzz, xi_rot, eta_rot, thermostat energy, barostat energy}"""
        out = np.zeros(12, np.float32)
        self._check(lib().ref_get_npt_state(self.h, _ptr(out)))
        return out

    def pressure_tensor(self, ts=None):
        """{Pxx, Pxy, Pxz, Pyy, Pyz, Pzz, P, KE} from the full-step ComputeThermo"""
        out = np.zeros(8, np.float32)
        self._check(lib().ref_get_pressure_tensor(self.h, self.timestep if ts is None else ts, _ptr(out)))
        return out

    def nvt_state(self):
        """Thermostat variables (xi, eta) of the NVT method."""
        out = np.zeros(2, np.float32)
        self._check(lib().ref_get_nvt_state(self.h, _ptr(out)))
        return out

    def nvt_reservoir_energy(self):
        out = np.zeros(1, np.float32)
        self._check(lib().ref_get_nvt_reservoir_energy(self.h, _ptr(out)))
        return float(out[0])

    def set_sorter(self, period):
        lib().ref_set_sorter(self.h, C.c_uint(period))

    def sort_now(self):
        self._check(lib().ref_sort_now(self.h))

    def set_timestep(self, ts):
        lib().ref_set_timestep(self.h, C.c_uint(ts))

    @property
    def timestep(self):
        return lib().ref_get_timestep(self.h)

    def compute_nlist(self, ts=0):
        self._check(lib().ref_compute_nlist(self.h, C.c_uint(ts)))

    def compute_forces(self, ts=0):
        self._check(lib().ref_compute_forces(self.h, C.c_uint(ts)))

    def run(self, nsteps):
        el = C.c_double(0)
        self._check(lib().ref_run(self.h, C.c_uint(nsteps), C.byref(el)))
        return el.value

    def step_one(self, ts):
        self._check(lib().ref_step_one(self.h, C.c_uint(ts)))

    def step_two(self, ts, net_force4):
        f = np.ascontiguousarray(net_force4, np.float32)
        self._check(lib().ref_step_two(self.h, C.c_uint(ts), _ptr(f)))

    def particles(self):
        N = self.N
        out = dict(pos=np.empty((N, 4), np.float32), vel=np.empty((N, 4), np.float32),
                   accel=np.empty((N, 3), np.float32), image=np.empty((N, 3), np.int32),
                   tag=np.empty(N, np.uint32), rtag=np.empty(N, np.uint32))
        lib().ref_get_particles(self.h, _ptr(out["pos"]), _ptr(out["vel"]), _ptr(out["accel"]), _ptr(out["image"]),
                                _ptr(out["tag"]), _ptr(out["rtag"]))
        return out

    def forces(self, virial=True):
        f = np.empty((self.N, 4), np.float32)
        v = np.empty((6, self.N), np.float32) if virial else None
        lib().ref_get_forces(self.h, _ptr(f), _ptr(v))
        return f, v

    def net_force(self):
        f = np.empty((self.N, 4), np.float32)
        lib().ref_get_net_force(self.h, _ptr(f))
        return f

    def nlist(self):
        """(n_neigh[N], head[N], nlist[alloc]) exactly as NeighborList::get*Array() hold them."""
        n = np.empty(self.N, np.uint32)
        head = np.empty(self.N, np.uint32)
        size = lib().ref_get_nlist_size(self.h)
        nl = np.empty(size, np.uint32)
        lib().ref_get_nlist(self.h, _ptr(n), _ptr(head), _ptr(nl), C.c_uint(size))
        return n, head, nl

    def nlist_stats(self):
        out = np.zeros(3, np.uint64)
        lib().ref_get_nlist_stats(self.h, _ptr(out))
        return dict(updates=int(out[0]), smallest_rebuild=int(out[1]))

    def cell_list(self):
        d = np.zeros(5, np.uint32)
        lib().ref_get_cell_dims(self.h, _ptr(d))
        dim, nmax, nadj = d[:3].copy(), int(d[3]), int(d[4])
        ncell = int(dim[0]) * int(dim[1]) * int(dim[2])
        size = np.empty(ncell, np.uint32)
        xyzf = np.empty((ncell, nmax, 4), np.float32)
        adj = np.empty((ncell, nadj), np.uint32)
        lib().ref_get_cell_list(self.h, _ptr(size), _ptr(xyzf), _ptr(adj))
        return dict(dim=dim, Nmax=nmax, cell_size=size, xyzf=xyzf, cell_adj=adj)

    def cell_tdb(self, ts=0):
        """switches CellList::setComputeTDB(true), recomputes the cell list at ts and returns the (ncell, Nmax, 4)
        (type, diameter, body, 0) entries; cell_list() afterwards returns the matching cell_size / xyzf"""
        self._check(lib().ref_cell_compute_tdb(self.h, C.c_uint(ts)))
        d = np.zeros(5, np.uint32)
        lib().ref_get_cell_dims(self.h, _ptr(d))
        tdb = np.empty((int(d[0]) * int(d[1]) * int(d[2]), int(d[3]), 4), np.float32)
        lib().ref_get_cell_tdb(self.h, _ptr(tdb))
        return tdb

    def thermo(self, ts=None):
        out = np.zeros(5, np.float64)
        self._check(lib().ref_get_thermo(self.h, C.c_uint(self.timestep if ts is None else ts), _ptr(out)))
        return dict(ke=out[0], pe=out[1], T=out[2], P=out[3], ndof=out[4])

    def lj_energy(self):
        return lib().ref_get_lj_energy(self.h)


def saru(seed1, seed2, seed3, n=3):
    """n draws of Saru(seed1,seed2,seed3).u32() and .s<float>(-1,1) from the reference's hoomd/Saru.h."""
    u = np.empty(n, np.uint32)
    f = np.empty(n, np.float32)
    lib().ref_saru(C.c_uint(seed1), C.c_uint(seed2), C.c_uint(seed3), C.c_uint(n), _ptr(u), _ptr(f))
    return u, f
