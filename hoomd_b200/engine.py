"""Thin object wrapper over the b200_sys_* engine C ABI (HOST numpy buffers in, HOST numpy buffers out).

This is the call a user of the library makes: the reference-facing job-script layer (hoomd_b200.md.*) drives this
class.  Method names follow the reference objects they stand for (hoomd/md/nlist.py, pair.py, integrate.py).
"""
import ctypes as C

import numpy as np

from . import _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class System:
    """SystemDefinition + nlist.cell + pair.lj + integrate.mode_standard + nve/langevin on one B200."""

    def __init__(self, pos4, box, ntypes=1, vel4=None, image=None, device=0, n_bond_types=0):
        L = _lib.lib()
        self.n_bond_types = int(n_bond_types)
        pos4 = np.ascontiguousarray(pos4, np.float32)
        self.N = int(pos4.shape[0])
        self.ntypes = int(ntypes)
        self.box = _lib.make_box(float(box[0]), float(box[1]), float(box[2]))
        self.h = C.c_void_p(L.b200_sys_create(self.N, C.byref(self.box), self.ntypes, int(device)))
        self._check(0)
        self.set_particles(pos4, vel4, image)

    def _check(self, rc):
        err = _lib.lib().b200_sys_last_error(self.h)
        if rc:
            raise RuntimeError(err.decode() if err else "hoomd_b200 error %d" % rc)
        if err and not err.startswith(b"conditional graph unavailable"):
            raise RuntimeError(err.decode())

    def close(self):
        if getattr(self, "h", None):
            _lib.lib().b200_sys_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- ParticleData -----------------------------------------------------------------------------------------
    def set_particles(self, pos4=None, vel4=None, image=None):
        a = None if pos4 is None else np.ascontiguousarray(pos4, np.float32)
        b = None if vel4 is None else np.ascontiguousarray(vel4, np.float32)
        c = None if image is None else np.ascontiguousarray(image, np.int32)
        self._check(_lib.lib().b200_sys_set_particles(self.h, _p(a), _p(b), _p(c)))

    def set_accel(self, accel3):
        a = np.ascontiguousarray(accel3, np.float32)
        self._check(_lib.lib().b200_sys_set_accel(self.h, _p(a)))

    def particles(self):
        N = self.N
        out = dict(pos=np.empty((N, 4), np.float32), vel=np.empty((N, 4), np.float32),
                   accel=np.empty((N, 3), np.float32), image=np.empty((N, 3), np.int32),
                   tag=np.empty(N, np.uint32), rtag=np.empty(N, np.uint32))
        self._check(_lib.lib().b200_sys_get_particles(self.h, _p(out["pos"]), _p(out["vel"]), _p(out["accel"]),
                                                      _p(out["image"]), _p(out["tag"]), _p(out["rtag"])))
        return out

    # ---- md.nlist.cell + md.pair.lj -------------------------------------------------------------------------
    def setup_lj(self, r_cut, epsilon=1.0, sigma=1.0, alpha=1.0, r_buff=0.4, check_period=1, shift_mode=0,
                 r_on=None):
        """lj1 = 4 eps sigma^12, lj2 = alpha 4 eps sigma^6 in double, then cast (hoomd/md/pair.py:602-609)."""
        nt = self.ntypes
        bc = lambda x: np.ascontiguousarray(np.broadcast_to(np.asarray(x, np.float64), (nt, nt)))
        eps, sig, alp, rc = bc(epsilon), bc(sigma), bc(alpha), bc(r_cut)
        lj1 = np.ascontiguousarray((4.0 * eps * sig ** 12).astype(np.float32))
        lj2 = np.ascontiguousarray((alp * 4.0 * eps * sig ** 6).astype(np.float32))
        rc = np.ascontiguousarray(rc.astype(np.float32))
        ron = None if r_on is None else np.ascontiguousarray(bc(r_on).astype(np.float32))
        self._check(_lib.lib().b200_sys_setup_lj(self.h, C.c_float(r_buff), int(check_period), _p(rc), _p(lj1),
                                                 _p(lj2), _p(ron), int(shift_mode)))
        self.lj1, self.lj2, self.rcut = lj1, lj2, rc

    def setup_pair(self, kind, r_cut, r_buff=0.4, check_period=1, shift_mode=0, r_on=None, **coeffs):
        """md.pair.<kind> in place of md.pair.lj: kind in gauss | yukawa | mie | force_shifted_lj | morse | lj, coefficients
        by their job-script names (epsilon, sigma, kappa, n, m, D0, alpha, r0), scalars or ntypes x ntypes arrays"""
        nt = self.ntypes
        bc = lambda x: np.ascontiguousarray(np.broadcast_to(np.asarray(x, np.float64), (nt, nt)))
        prm = _lib.pair_params(kind, nt, **coeffs)
        rc = np.ascontiguousarray(bc(r_cut).astype(np.float32))
        ron = None if r_on is None else np.ascontiguousarray(bc(r_on).astype(np.float32))
        self._check(_lib.lib().b200_sys_setup_pair(self.h, _lib.PAIR_KINDS[kind], C.c_float(r_buff), int(check_period),
                                                   _p(rc), _p(prm), _p(ron), int(shift_mode)))
        self.pair_params, self.rcut = prm, rc

    # ---- BondData + md.bond.* + nlist exclusions -------------------------------------------------------------
    def set_bonds(self, bonds, types=None, n_bond_types=None):
        """bonds: (nbonds, 2) particle TAGS; types: bond type per bond (default 0)"""
        b = np.ascontiguousarray(bonds, np.uint32).reshape(-1, 2)
        t = np.zeros(len(b), np.uint32) if types is None else np.ascontiguousarray(types, np.uint32)
        self.n_bond_types = int(n_bond_types or self.n_bond_types or 1)
        self._check(_lib.lib().b200_sys_set_bonds(self.h, len(b), _p(b), _p(t), self.n_bond_types))

    def setup_fene(self, k=30.0, r0=1.5, epsilon=1.0, sigma=1.0):
        """md.bond.fene + bond_coeff.set: lj1 = 4 eps sigma^12, lj2 = 4 eps sigma^6 (hoomd/md/bond.py:367-372)"""
        nb = self.n_bond_types
        bc = lambda x: np.broadcast_to(np.asarray(x, np.float64), (nb,))
        K = np.ascontiguousarray(bc(k), np.float32)
        R = np.ascontiguousarray(bc(r0), np.float32)
        lj1 = np.ascontiguousarray(4.0 * bc(epsilon) * bc(sigma) ** 12, np.float32)
        lj2 = np.ascontiguousarray(4.0 * bc(epsilon) * bc(sigma) ** 6, np.float32)
        self._check(_lib.lib().b200_sys_setup_bond_force(self.h, 0, _p(K), _p(R), _p(lj1), _p(lj2)))
        self.bond_params = (K, R, lj1, lj2)

    def setup_harmonic(self, k, r0):
        """md.bond.harmonic + bond_coeff.set (hoomd/md/bond.py:290-296)"""
        nb = self.n_bond_types
        K = np.ascontiguousarray(np.broadcast_to(np.asarray(k, np.float64), (nb,)), np.float32)
        R = np.ascontiguousarray(np.broadcast_to(np.asarray(r0, np.float64), (nb,)), np.float32)
        self._check(_lib.lib().b200_sys_setup_bond_force(self.h, 1, _p(K), _p(R), None, None))
        self.bond_params = (K, R)

    def exclusions_from_bonds(self, on=True):
        """nlist.reset_exclusions(['bond']) / reset_exclusions([])"""
        self._check(_lib.lib().b200_sys_nlist_exclusions_from_bonds(self.h, int(on)))

    def force_component(self, which, virial=False):
        """per-ForceCompute arrays after compute_forces: 0 = pair, 1 = bond"""
        f = np.empty((self.N, 4), np.float32)
        v = np.empty((6, self.N), np.float32) if virial else None
        self._check(_lib.lib().b200_sys_get_force_component(self.h, int(which), _p(f), _p(v)))
        return f, v

    def set_virial(self, on=True):
        self._check(_lib.lib().b200_sys_set_virial(self.h, int(on)))

    # ---- md.integrate.* -------------------------------------------------------------------------------------
    def setup_integrator(self, dt):
        self._check(_lib.lib().b200_sys_setup_integrator(self.h, C.c_float(dt)))

    def add_nve(self, limit=None, zero_force=False):
        self._check(_lib.lib().b200_sys_add_nve(self.h, int(limit is not None), C.c_float(limit or 0.0),
                                                int(zero_force)))

    def add_langevin(self, kT, seed, gamma=None, noiseless_t=False, dscale=None, tally=False):
        """md.integrate.langevin(group, kT, seed, dscale, tally, noiseless_t); dscale = lambda: gamma = lambda * diameter"""
        g = None if gamma is None else np.ascontiguousarray(np.broadcast_to(gamma, (self.ntypes,)), np.float32)
        self._check(_lib.lib().b200_sys_add_langevin(self.h, C.c_float(kT), C.c_uint(seed), _p(g), int(noiseless_t)))
        if dscale is not None or tally:
            self._check(_lib.lib().b200_sys_set_langevin_options(self.h, int(dscale is not None),
                                                                 C.c_float(dscale or 0.0), int(bool(tally))))

    add_langevin_ex = add_langevin  # the name oracle.ref.RefSystem uses for the extended form

    def langevin_reservoir_energy(self):
        out = C.c_double(0)
        self._check(_lib.lib().b200_sys_get_langevin_reservoir_energy(self.h, C.byref(out)))
        return out.value

    def set_diameters(self, diameter_by_tag):
        d = np.ascontiguousarray(diameter_by_tag, np.float32)
        assert len(d) == self.N
        self._check(_lib.lib().b200_sys_set_diameters(self.h, _p(d)))

    def set_group(self, tags=None):
        """the ParticleGroup of the integration method: member tags (hoomd.group.tag_list / tags / type), None = group.all()"""
        t = np.zeros(0, np.uint32) if tags is None else np.ascontiguousarray(tags, np.uint32)
        self._check(_lib.lib().b200_sys_set_group(self.h, _p(t) if len(t) else None, len(t)))

    def set_box(self, Lx, Ly, Lz, scale_particles=True):
        """hoomd.update.box_resize at a period boundary: one BoxResizeUpdater::update towards the box (Lx, Ly, Lz)"""
        self._check(_lib.lib().b200_sys_set_box(self.h, C.c_float(Lx), C.c_float(Ly), C.c_float(Lz), int(scale_particles)))
        self.box = _lib.make_box(float(Lx), float(Ly), float(Lz))

    def box_lengths(self):
        out = np.zeros(3, np.float32)
        _lib.lib().b200_sys_get_box(self.h, _p(out))
        return out

    def set_kT(self, kT):
        """set point of the active langevin / nvt method (setT); a variant is evaluated by the caller, once per step"""
        self._check(_lib.lib().b200_sys_set_temperature(self.h, C.c_float(kT)))
        if hasattr(self, "_nvt"):
            self._nvt = (np.float32(kT), self._nvt[1])

    def set_dist_check(self, on=True):
        """md.nlist.cell(dist_check=...): off = rebuild whenever the check period is hit"""
        self._check(_lib.lib().b200_sys_set_dist_check(self.h, int(on)))

    def add_nvt(self, kT, tau):
        """md.integrate.nvt(group=all, kT, tau): TwoStepNVTMTK (hoomd/md/integrate.py:163-260)"""
        self._check(_lib.lib().b200_sys_add_nvt(self.h, C.c_float(kT), C.c_float(tau)))
        self._nvt = (np.float32(kT), np.float32(tau))

    def nvt_reservoir_energy(self):
        """log quantity nvt_mtk_reservoir_energy: g kT (xi^2 tau^2 / 2 + eta), evaluated like TwoStepNVTMTK::getLogValue
        (hoomd/md/TwoStepNVTMTK.cc:82-107: FP32 bracket, product with the double set point)"""
        xi, eta = self.nvt_state()
        kT, tau = self._nvt
        g = np.float32(3 * self.N - 3)
        inner = np.float32(np.float32(np.float32(np.float32(xi * xi) * tau) * tau) / np.float32(2.0)) + eta
        return float(np.float32(np.float64(g) * np.float64(kT) * np.float64(np.float32(inner))))

    def set_nvt_state(self, xi, eta):
        """restore the thermostat variables of an earlier run (restart / engine re-configuration)"""
        st = np.array([xi, eta], np.float32)
        self._check(_lib.lib().b200_sys_set_nvt_state(self.h, _p(st)))

    def nvt_state(self):
        """(xi, eta) of the Nose-Hoover thermostat"""
        out = np.zeros(2, np.float32)
        self._check(_lib.lib().b200_sys_get_nvt_state(self.h, _p(out)))
        return out

    def set_dpd_thermostat(self, kT, seed):
        """md.pair.dpd(kT, seed): the DPD pair force set up with setup_pair("dpd_conservative", A=..., gamma=...) becomes the
        pairwise thermostat (drag + random forces per pair); kT <= 0 switches back to the conservative force"""
        self._check(_lib.lib().b200_sys_set_dpd_thermostat(self.h, C.c_float(kT), C.c_uint(seed)))

    COUPLE = {"none": 0, "xy": 1, "xz": 2, "yz": 3, "xyz": 4}

    def add_npt(self, kT, tau, P, tauP, couple="xyz", x=True, y=True, z=True, nph=False, rescale_all=False, gamma=0.0):
        """md.integrate.npt(group, kT, tau, P, tauP, couple, x, y, z, nph, rescale_all, gamma): TwoStepNPTMTK on an
        orthorhombic box (hoomd/md/integrate.py:406-506).  Steps are host-driven: the box changes every step."""
        flags = (1 if x else 0) | (2 if y else 0) | (4 if z else 0)
        self._check(_lib.lib().b200_sys_add_npt(self.h, C.c_float(kT), C.c_float(tau), C.c_float(P), C.c_float(tauP),
                                                self.COUPLE[couple], flags, int(nph), int(rescale_all), C.c_float(gamma)))

    def npt_state(self):
        """{eta, xi, nuxx, nuxy, nuxz, nuyy, nuyz, nuzz, xi_rot, eta_rot, npt_thermostat_energy, npt_barostat_energy}"""
        out = np.zeros(12, np.float32)
        self._check(_lib.lib().b200_sys_get_npt_state(self.h, _p(out)))
        return out

    def set_npt_state(self, vars10):
        st = np.ascontiguousarray(vars10, np.float32)
        assert st.size >= 10
        self._check(_lib.lib().b200_sys_set_npt_state(self.h, _p(st)))

    def pressure_tensor(self):
        """{Pxx, Pxy, Pxz, Pyy, Pyz, Pzz, P, KE} of the integration group at the current step (needs set_virial)"""
        out = np.zeros(8, np.float32)
        self._check(_lib.lib().b200_sys_get_pressure_tensor(self.h, _p(out)))
        return out

    def set_sorter(self, period):
        self._check(_lib.lib().b200_sys_set_sorter(self.h, int(period)))

    def sort_now(self):
        self._check(_lib.lib().b200_sys_sort_now(self.h))

    def set_tuning(self, pair_tpp=0, use_graph=2):
        self._check(_lib.lib().b200_sys_set_tuning(self.h, int(pair_tpp), int(use_graph)))

    def set_pair_mode(self, mode=-1, tile=(0, 0, 0), tpp=0):
        """Pair-force path of the step loop: 0 = walk the reference-layout neighbor list, 1 = shared-memory cell tiles
        (particles kept in cell order), -1 = tiles when supported.  tile = cells per CTA, tpp = lanes per particle."""
        self._check(_lib.lib().b200_sys_set_pair_mode(self.h, int(mode), int(tile[0]), int(tile[1]), int(tile[2]),
                                                      int(tpp)))

    def pair_mode(self):
        out = (C.c_uint * 8)()
        _lib.lib().b200_sys_get_pair_mode(self.h, out)
        return dict(tile=bool(out[0]), cells_per_tile=(int(out[1]), int(out[2]), int(out[3])), cap=int(out[4]),
                    tpp=int(out[5]), row_capacity=int(out[6]), tiles=int(out[7]))

    def stepping(self):
        """how the last run stepped: use_graph mode, MD steps per graph unit (0 = data dependent), multi-step kernel on/off,
        graph units launched by the last run"""
        out = (C.c_uint * 4)()
        _lib.lib().b200_sys_get_stepping(self.h, out)
        return dict(use_graph=int(out[0]), steps_per_unit=int(out[1]), multistep=bool(out[2]), units_last_run=int(out[3]))

    def set_multistep(self, on=True):
        """several MD steps per launch (k_md_steps) where it applies (default on); off = the one-step kernel everywhere"""
        self._check(_lib.lib().b200_sys_set_multistep(self.h, int(bool(on))))

    def set_timestep(self, ts):
        _lib.lib().b200_sys_set_timestep(self.h, int(ts))

    @property
    def timestep(self):
        return int(_lib.lib().b200_sys_get_timestep(self.h))

    # ---- compute / run -----------------------------------------------------------------------------------------
    def compute_nlist(self, ts=0):
        self._check(_lib.lib().b200_sys_compute_nlist(self.h, int(ts)))

    def compute_forces(self, ts=0):
        self._check(_lib.lib().b200_sys_compute_forces(self.h, int(ts)))

    def run(self, nsteps):
        """System::run(nsteps).  Returns (wall seconds of the step loop, CUDA-event milliseconds)."""
        el, ms = C.c_double(0), C.c_double(0)
        self._check(_lib.lib().b200_sys_run(self.h, int(nsteps), C.byref(el), C.byref(ms)))
        return el.value, ms.value

    def step_one(self, ts):
        self._check(_lib.lib().b200_sys_step_one(self.h, int(ts)))

    def step_two(self, ts, net_force4=None):
        f = None if net_force4 is None else np.ascontiguousarray(net_force4, np.float32)
        self._check(_lib.lib().b200_sys_step_two(self.h, int(ts), _p(f)))

    def forces(self, virial=True):
        f = np.empty((self.N, 4), np.float32)
        v = np.empty((6, self.N), np.float32) if virial else None
        self._check(_lib.lib().b200_sys_get_forces(self.h, _p(f), _p(v)))
        return f, v

    net_force = lambda self: self.forces(False)[0]

    def nlist(self, counts_only=False):
        n = np.empty(self.N, np.uint32)
        head = np.empty(self.N, np.uint32)
        if counts_only:
            self._check(_lib.lib().b200_sys_get_nlist(self.h, _p(n), _p(head), None, 0))
            return n, head, None
        size = int(_lib.lib().b200_sys_get_nlist_size(self.h))
        nl = np.empty(size, np.uint32)
        self._check(_lib.lib().b200_sys_get_nlist(self.h, _p(n), _p(head), _p(nl), size))
        return n, head, nl

    def stats(self):
        out = (C.c_ulonglong * 8)()
        _lib.lib().b200_sys_get_stats(self.h, out)
        return dict(updates=int(out[0]), forced=int(out[1]), overflows=int(out[2]), kernels_last_run=int(out[3]),
                    dim=(int(out[4]), int(out[5]), int(out[6])), Nmax0=int(out[7]))

    def time_kernel(self, which, iters=20):
        """average CUDA-event milliseconds per launch: which=0 pair kernel, 1 cell list + nlist rebuild, 2 the single-kernel
        MD step (runs on a scratch copy of the integrator state; the system is not advanced)"""
        ms = C.c_double(0)
        self._check(_lib.lib().b200_sys_time_kernel(self.h, int(which), int(iters), C.byref(ms)))
        return ms.value

    def thermo(self):
        out = np.zeros(5, np.float64)
        self._check(_lib.lib().b200_sys_get_thermo(self.h, out.ctypes.data_as(C.POINTER(C.c_double))))
        return dict(ke=out[0], pe=out[1], T=out[2], P=out[3], ndof=out[4])
