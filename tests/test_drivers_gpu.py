"""Level-1 boundary tests: every kernel driver of include/hoomd_b200.h called the way the reference's `...GPU.cc` classes
would call it -- raw device pointers into CALLER-OWNED buffers (torch tensors here, GPUArray there), legacy stream 0, no
engine object -- and compared with the UNMODIFIED reference CPU classes (oracle/_ref) on the same inputs.

  cell list      b200_cell_list_dimensions + b200_compute_cell_list + b200_cell_list_export_padded   vs CellList's
                 getCellSizeArray / getXYZFArray / getTDBArray (hoomd/CellList.h:263-300);
                 b200_cell_list_build_adjacency vs getCellAdjArray (hoomd/CellList.cc:373-436)
  chain          b200_nlist_build_head_list -> b200_compute_nlist_binned -> b200_compute_lj_forces ->
                 b200_integrator_sum_net_force -> b200_nve_step_two (+ b200_nve_step_one) on raw pointers vs RefSystem
  distance check b200_nlist_needs_update_check vs NeighborList::distanceCheck semantics (hoomd/md/NeighborList.cc:964-1040)
  sparse group   b200_nve_step_one / _two and b200_langevin_step_two (+ dscale, tally) over every third particle vs the
                 reference's group integration (hoomd/md/TwoStepNVEGPU.cc:48-70: m_group->getIndexArray())
  sorter         b200_sfc_generate_sorted_order + b200_sfc_apply_sorted_order vs SFCPackUpdater
"""
import ctypes as C

import numpy as np
import pytest

import golden_util as G

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref():
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref/libhoomd_ref.so not built")
    return R


@pytest.fixture(scope="module")
def env():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import hoomd_b200
    from hoomd_b200 import _lib
    return hoomd_b200, _lib, _lib.lib(), torch


def _dev(torch, a):
    a = np.ascontiguousarray(a)
    if a.dtype == np.uint32:
        a = a.view(np.int32)
    return torch.from_numpy(a).cuda()


def _u32(t):
    return t.cpu().numpy().view(np.uint32)


def _cell_list(env, pos, L, r_list, ntypes):
    """b200_cell_list_dimensions + b200_compute_cell_list into torch buffers; returns the argument block and the buffers"""
    hb, _lib, Lb, torch = env
    N = len(pos)
    box = _lib.make_box(L, L, L)
    dim = _lib.Dim3()
    gw = (C.c_float * 3)()
    Lb.b200_cell_list_dimensions(C.byref(box), C.c_float(r_list), 1, C.c_float(0.0), C.byref(dim), gw)
    ncells = dim.x * dim.y * dim.z
    z = lambda n, dt=torch.int32: torch.zeros(max(n, 1), dtype=dt, device="cuda")
    buf = dict(pos=_dev(torch, pos), cell_start=z(ncells + 1), cell_xyzf=torch.zeros((N, 4), dtype=torch.float32, device="cuda"),
               cell_type=z(N), cell_idx=z(N), cond=z(3), cell_count=z(ncells), pcell=z(N), pslot=z(N), scratch=z(N))
    a = _lib.CellListArgs()
    a.d_cell_start, a.d_cell_xyzf = buf["cell_start"].data_ptr(), buf["cell_xyzf"].data_ptr()
    a.d_cell_type = buf["cell_type"].data_ptr() if ntypes > 1 else None
    a.d_cell_idx, a.d_conditions = buf["cell_idx"].data_ptr(), buf["cond"].data_ptr()
    a.d_cell_count, a.d_particle_cell = buf["cell_count"].data_ptr(), buf["pcell"].data_ptr()
    a.d_particle_slot, a.d_scratch_idx = buf["pslot"].data_ptr(), buf["scratch"].data_ptr()
    a.d_pos, a.d_charge, a.N, a.n_ghost, a.flag_charge, a.flag_type = buf["pos"].data_ptr(), None, N, 0, 0, 0
    a.box, a.dim = box, dim
    for d in range(3):
        a.ghost_width[d] = gw[d]
    assert Lb.b200_compute_cell_list(C.byref(a), None) == 0
    torch.cuda.synchronize()
    return a, buf, box, dim


@pytest.mark.parametrize("N,L,ntypes,seed", [(3000, 14.0, 2, 7), (500, 8.0, 1, 3), (20000, 30.0, 1, 5)])
def test_cell_list_drivers_vs_reference_arrays(ref, env, N, L, ntypes, seed):
    hb, _lib, Lb, torch = env
    pos = hb.workloads.random_system(N, L, seed=seed, ntypes=ntypes)
    r = ref.RefSystem(pos, (L, L, L), ntypes=ntypes)
    r.setup_lj(r_cut=3.0, r_buff=0.4, full=True)
    r.compute_nlist(0)
    tdb_ref = r.cell_tdb(0)                      # switches TDB on and recomputes: what NeighborListGPUBinned asks for
    cl = r.cell_list()
    a, buf, box, dim = _cell_list(env, pos, L, 3.4, ntypes)
    assert (dim.x, dim.y, dim.z) == tuple(int(v) for v in cl["dim"])
    ncells, Nmax = dim.x * dim.y * dim.z, cl["Nmax"]
    # conditions (hoomd/CellList.cc:575-621): .x = largest cell occupancy seen, .y / .z = NaN / out-of-box particle + 1
    cond = _u32(buf["cond"])
    assert cond[0] == cl["cell_size"].max() and cond[1] == 0 and cond[2] == 0
    # the contiguous layout itself: an exclusive scan of the reference's cell sizes
    start = _u32(buf["cell_start"])
    assert np.array_equal(np.diff(start.astype(np.int64)), cl["cell_size"]) and start[-1] == N
    # export into the reference's padded arrays
    d_size = torch.zeros(ncells, dtype=torch.int32, device="cuda")
    d_xyzf = torch.zeros((ncells, Nmax, 4), dtype=torch.float32, device="cuda")
    d_tdb = torch.zeros((ncells, Nmax, 4), dtype=torch.float32, device="cuda")
    d_cond = torch.zeros(3, dtype=torch.int32, device="cuda")
    assert Lb.b200_cell_list_export_padded(d_size.data_ptr(), d_xyzf.data_ptr(), d_tdb.data_ptr(), d_cond.data_ptr(),
                                           buf["cell_start"].data_ptr(), buf["cell_xyzf"].data_ptr(),
                                           buf["cell_idx"].data_ptr(), buf["pos"].data_ptr(), None, None, ncells, Nmax,
                                           None) == 0
    torch.cuda.synchronize()
    size = _u32(d_size)
    assert np.array_equal(size, cl["cell_size"])
    assert _u32(d_cond)[0] == 0                   # nothing exceeded the reference's Nmax
    valid = np.arange(Nmax)[None, :] < size[:, None]
    xyzf, tdb = d_xyzf.cpu().numpy(), d_tdb.cpu().numpy()
    # bit-equal entries in the reference's order (ascending particle index inside a cell, hoomd/CellList.cc:474-563)
    assert np.array_equal(xyzf.view(np.uint32)[valid], cl["xyzf"].view(np.uint32)[valid])
    assert np.array_equal(tdb.view(np.uint32)[valid], tdb_ref.view(np.uint32)[valid])
    # overflow protocol: an Nmax that is too small drops the surplus and records the needed Nmax (hoomd/CellList.cc:575-621)
    small = max(1, int(size.max()) - 1)
    d_cond.zero_()
    d_x2 = torch.zeros((ncells, small, 4), dtype=torch.float32, device="cuda")
    assert Lb.b200_cell_list_export_padded(d_size.data_ptr(), d_x2.data_ptr(), None, d_cond.data_ptr(),
                                           buf["cell_start"].data_ptr(), buf["cell_xyzf"].data_ptr(), None,
                                           buf["pos"].data_ptr(), None, None, ncells, small, None) == 0
    torch.cuda.synchronize()
    assert _u32(d_cond)[0] == size.max()
    # adjacency (CellList::initializeCellAdj): sorted, de-duplicated wrapped neighbor cells
    nadj = cl["cell_adj"].shape[1]
    d_adj = torch.zeros((ncells, nadj), dtype=torch.int32, device="cuda")
    assert Lb.b200_cell_list_build_adjacency(d_adj.data_ptr(), dim, 1, nadj, None) == 0
    torch.cuda.synchronize()
    assert np.array_equal(_u32(d_adj), cl["cell_adj"])


def test_driver_chain_on_raw_pointers_vs_reference(ref, env):
    """head list -> binned build -> LJ force -> net-force sum -> NVE step two (and step one), all on caller buffers"""
    hb, _lib, Lb, torch = env
    N, L, ntypes = 4000, 16.0, 2
    rng = np.random.default_rng(21)
    pos = hb.workloads.random_system(N, L, seed=9, ntypes=ntypes)
    # keep random overlaps soft: small epsilon
    lj = dict(r_cut=[[2.5, 2.0], [2.0, 2.2]], epsilon=[[1e-3, 1.5e-3], [1.5e-3, 5e-4]], sigma=[[1.0, 0.8], [0.8, 0.88]],
              r_buff=0.3, shift_mode=1)
    vel = np.zeros((N, 4), np.float32)
    vel[:, :3] = rng.normal(0, 1.0, (N, 3))
    vel[:, 3] = rng.uniform(0.5, 2.0, N)
    r = ref.RefSystem(pos, (L, L, L), ntypes=ntypes, vel4=vel)
    r.setup_lj(full=True, **lj)
    r.set_virial(True)
    r.compute_forces(0)
    rn, rh, rl = r.nlist()
    fr, vr = r.forces()
    # Nmax[type] as the reference grew it: the head list is the exclusive scan of Nmax[type_i] (NeighborList.cc:1336-1357)
    types = pos[:, 3].view(np.int32)
    Nmax = np.zeros(ntypes, np.uint32)
    for t in range(ntypes):
        i = int(np.nonzero(types[:-1] == t)[0][0])
        Nmax[t] = rh[i + 1] - rh[i]
    # tables as NeighborList::updateRList / PotentialPair::setRcut hold them (FP32)
    rc = np.asarray(lj["r_cut"], np.float32)
    rlistsq = ((rc + np.float32(lj["r_buff"])) ** 2).astype(np.float32)
    eps, sig = np.asarray(lj["epsilon"], np.float64), np.asarray(lj["sigma"], np.float64)
    params = np.stack([(4.0 * eps * sig ** 12), (4.0 * eps * sig ** 6)], axis=-1).astype(np.float32)
    a, buf, box, dim = _cell_list(env, pos, L, float(rc.max()) + lj["r_buff"], ntypes)
    d_Nmax, d_head, d_req = _dev(torch, Nmax), torch.zeros(N, dtype=torch.int32, device="cuda"), torch.zeros(1, dtype=torch.int32, device="cuda")
    assert Lb.b200_nlist_build_head_list(d_head.data_ptr(), d_req.data_ptr(), d_Nmax.data_ptr(), buf["pos"].data_ptr(), N,
                                         ntypes, None) == 0
    torch.cuda.synchronize()
    assert np.array_equal(_u32(d_head), rh)
    req = int(_u32(d_req)[0])
    d_nlist = torch.zeros(req, dtype=torch.int32, device="cuda")
    d_nn = torch.zeros(N, dtype=torch.int32, device="cuda")
    d_last = torch.zeros((N, 4), dtype=torch.float32, device="cuda")
    d_cond = torch.zeros(ntypes, dtype=torch.int32, device="cuda")
    d_rc, d_rl = _dev(torch, rc), _dev(torch, rlistsq)
    na = _lib.NlistArgs()
    na.d_nlist, na.d_n_neigh, na.d_last_pos, na.d_conditions = d_nlist.data_ptr(), d_nn.data_ptr(), d_last.data_ptr(), d_cond.data_ptr()
    na.d_Nmax, na.d_head_list, na.d_pos, na.N = d_Nmax.data_ptr(), d_head.data_ptr(), buf["pos"].data_ptr(), N
    na.d_cell_start, na.d_cell_xyzf, na.d_cell_type = buf["cell_start"].data_ptr(), buf["cell_xyzf"].data_ptr(), buf["cell_type"].data_ptr()
    na.dim, na.box, na.d_r_cut, na.d_r_listsq, na.ntypes, na.half_storage, na.cand_hint = dim, box, d_rc.data_ptr(), d_rl.data_ptr(), ntypes, 0, 0
    assert Lb.b200_compute_nlist_binned(C.byref(na), None) == 0
    torch.cuda.synchronize()
    gn, gl = _u32(d_nn), _u32(d_nlist)
    assert np.array_equal(gn, rn) and (_u32(d_cond) <= Nmax).all()
    for i in range(N):
        assert np.array_equal(gl[rh[i]:rh[i] + rn[i]], rl[rh[i]:rh[i] + rn[i]]), i
    assert np.array_equal(d_last.cpu().numpy()[:, :3].view(np.uint32), pos[:, :3].view(np.uint32))   # setLastUpdatedPos
    # pair force into a caller force / virial array with the reference's pitch rule
    pitch = N + (16 - (N & 15))
    d_f = torch.zeros((N, 4), dtype=torch.float32, device="cuda")
    d_v = torch.zeros((6, pitch), dtype=torch.float32, device="cuda")
    pa = _lib.PairArgs()
    pa.d_force, pa.d_virial, pa.virial_pitch, pa.N, pa.d_pos, pa.box = d_f.data_ptr(), d_v.data_ptr(), pitch, N, buf["pos"].data_ptr(), box
    pa.d_n_neigh, pa.d_nlist, pa.d_head_list = d_nn.data_ptr(), d_nlist.data_ptr(), d_head.data_ptr()
    d_rcsq, d_prm = _dev(torch, (rc * rc).astype(np.float32)), _dev(torch, params)
    pa.d_rcutsq, pa.d_ronsq, pa.ntypes, pa.shift_mode, pa.compute_virial = d_rcsq.data_ptr(), d_rcsq.data_ptr(), ntypes, 1, 1
    pa.threads_per_particle, pa.block_size, pa.wrap_skip_margin = 0, 0, 0.0
    assert Lb.b200_compute_lj_forces(C.byref(pa), d_prm.data_ptr(), None) == 0
    torch.cuda.synchronize()
    fg, vg = d_f.cpu().numpy(), d_v.cpu().numpy()[:, :N]
    sc = G.abs_scales(lambda o: o.setup_lj(**lj), pos, (L, L, L), ntypes=ntypes)
    G.assert_parity(fr, vr, fg, vg, sc, "b200_compute_lj_forces on caller buffers")
    # Integrator::computeNetForceGPU: net = clear + force list (a second, synthetic force array stands for another ForceCompute)
    f2 = rng.normal(0, 0.01, (N, 4)).astype(np.float32)
    v2 = rng.normal(0, 0.01, (6, pitch)).astype(np.float32)
    d_f2, d_v2 = _dev(torch, f2), _dev(torch, v2)
    d_net = torch.full((N, 4), 7.0, dtype=torch.float32, device="cuda")
    d_netv = torch.full((6, pitch), 7.0, dtype=torch.float32, device="cuda")
    fl = _lib.ForceList()
    fl.f[0], fl.f[1], fl.v[0], fl.v[1], fl.vpitch[0], fl.vpitch[1] = d_f.data_ptr(), d_f2.data_ptr(), d_v.data_ptr(), d_v2.data_ptr(), pitch, pitch
    assert Lb.b200_integrator_sum_net_force(d_net.data_ptr(), d_netv.data_ptr(), pitch, C.byref(fl), N, 1, 1, None) == 0
    torch.cuda.synchronize()
    assert np.array_equal(d_net.cpu().numpy().view(np.uint32), (fg + f2).view(np.uint32))
    assert np.array_equal(d_netv.cpu().numpy()[:, :N].view(np.uint32), (vg + v2[:, :N]).view(np.uint32))
    # without clear: accumulates onto what is there (hoomd/Integrator.cu:57-113)
    assert Lb.b200_integrator_sum_net_force(d_net.data_ptr(), d_netv.data_ptr(), pitch, C.byref(fl), N, 0, 1, None) == 0
    torch.cuda.synchronize()
    assert np.array_equal(d_net.cpu().numpy().view(np.uint32), ((fg + f2) + fg + f2).view(np.uint32))
    # NVE half steps on caller buffers, bit for bit against TwoStepNVE on the same net force
    net = (fg + f2).astype(np.float32)
    accel = rng.normal(0, 5.0, (N, 3)).astype(np.float32)
    for s_ in (r,):
        s_.setup_integrator(0.005)
        s_.add_nve()
        s_.set_accel(accel)
    d_vel, d_acc, d_img = _dev(torch, vel), _dev(torch, accel), torch.zeros((N, 3), dtype=torch.int32, device="cuda")
    d_pos2 = _dev(torch, pos)
    assert Lb.b200_nve_step_one(d_pos2.data_ptr(), d_vel.data_ptr(), d_acc.data_ptr(), d_img.data_ptr(), None, N, C.byref(box),
                                C.c_float(0.005), 0, C.c_float(0.0), 0, None) == 0
    r.step_one(0)
    d_netf = _dev(torch, net)
    assert Lb.b200_nve_step_two(d_vel.data_ptr(), d_acc.data_ptr(), None, N, d_netf.data_ptr(), C.c_float(0.005), 0,
                                C.c_float(0.0), 0, None) == 0
    r.step_two(0, net)
    torch.cuda.synchronize()
    p = r.particles()
    for name, t in (("pos", d_pos2), ("vel", d_vel), ("accel", d_acc), ("image", d_img)):
        assert np.array_equal(t.cpu().numpy().view(np.uint32), p[name].view(np.uint32)), name


def test_distance_check_driver(env):
    """NeighborList::distanceCheck: result = checkn iff some particle moved >= ((rcut_max + r_buff) lambda_min - rcut_max) / 2
    from lambda * last_pos (minimum image); untouched otherwise"""
    hb, _lib, Lb, torch = env
    N, L, r_buff = 5000, 12.0, 0.4
    pos = hb.workloads.random_system(N, L, seed=2, ntypes=2)
    box = _lib.make_box(L, L, L)
    rcut_max = np.array([3.0, 2.0], np.float32)
    d_rm = _dev(torch, rcut_max)
    lam = (C.c_float * 3)(1.0, 1.0, 1.0)
    d_last = _dev(torch, pos)

    def check(p, lam3=lam, lmin=1.0, checkn=17):
        d_res = torch.zeros(1, dtype=torch.int32, device="cuda")
        d_p = _dev(torch, p)
        assert Lb.b200_nlist_needs_update_check(d_res.data_ptr(), d_last.data_ptr(), d_p.data_ptr(), N, C.byref(box),
                                                d_rm.data_ptr(), C.c_float(r_buff), 2, C.c_float(lmin), lam3, checkn, None) == 0
        torch.cuda.synchronize()
        return int(_u32(d_res)[0])

    assert check(pos) == 0
    moved = pos.copy()
    moved[:, 0] += np.float32(0.19)                       # everyone just under r_buff / 2
    moved[:, 0] = np.where(moved[:, 0] >= L / 2, moved[:, 0] - np.float32(L), moved[:, 0])   # wrapped: minimum image applies
    assert check(moved) == 0
    one = moved.copy()
    one[1234, 1] += np.float32(0.08)                      # sqrt(0.19^2 + 0.08^2) = 0.206 >= 0.2
    assert check(one) == 17
    # a compressed box: lambda scales last_pos, the threshold shrinks to ((rcut_max_i + r_buff) * lambda_min - rcut_max_i) / 2
    lam2, lmin = (C.c_float * 3)(0.99, 0.99, 0.99), 0.99
    scaled = pos.copy()
    scaled[:, :3] *= np.float32(0.99)
    assert check(scaled, lam2, lmin) == 0                 # affine motion alone never triggers
    t0 = ((np.float32(3.0) + np.float32(r_buff)) * np.float32(lmin) - np.float32(3.0)) / np.float32(2.0)   # type 0: 0.183
    k = int(np.nonzero(pos[:, 3].view(np.int32) == 0)[0][5])
    kick = scaled.copy()
    kick[k, 2] += np.float32(t0 * 1.01)
    if kick[k, 2] >= L / 2:
        kick[k, 2] -= np.float32(2 * t0 * 1.01)
    assert check(kick, lam2, lmin, checkn=3) == 3
    kick[k, 2] = scaled[k, 2] + np.float32(t0 * 0.9) * (1 if scaled[k, 2] < 0 else -1)
    assert check(kick, lam2, lmin, checkn=3) == 0


def test_sparse_group_integration_vs_reference(ref, env):
    """every third particle through the NVE and Langevin drivers (d_group_members = the group's index array): members
    bit-identical to the reference's group integration, non-members untouched; Langevin with dscale (gamma = lambda *
    diameter) and the reservoir-energy tally (hoomd/md/TwoStepLangevin.cc:287-318,416-417)"""
    hb, _lib, Lb, torch = env
    N, L, dt = 3000, 11.0, 0.005
    rng = np.random.default_rng(8)
    pos = hb.workloads.random_system(N, L, seed=6, ntypes=2)
    vel = np.zeros((N, 4), np.float32)
    vel[:, :3] = rng.normal(0, 1.2, (N, 3))
    vel[:, 3] = rng.uniform(0.5, 2.0, N)
    accel = rng.normal(0, 8.0, (N, 3)).astype(np.float32)
    net = rng.normal(0, 15.0, (N, 4)).astype(np.float32)
    diam = rng.uniform(0.8, 1.6, N).astype(np.float32)
    members = np.arange(1, N, 3, dtype=np.uint32)
    box = _lib.make_box(L, L, L)
    d_grp = _dev(torch, members)
    mask = np.zeros(N, bool)
    mask[members] = True

    def reference(method, **kw):
        r = ref.RefSystem(pos, (L, L, L), ntypes=2, vel4=vel)
        r.setup_lj(r_cut=2.0, r_buff=0.3, full=True)
        r.set_diameters(diam)
        r.setup_integrator(dt)
        r.set_group(members)
        if method == "nve":
            r.add_nve()
        else:
            r.add_langevin_ex(1.3, 11, **kw)
        r.set_accel(accel)
        return r

    # ---- NVE ----
    r = reference("nve")
    r.step_one(5)
    r.step_two(5, net)
    p = r.particles()
    d_pos, d_vel, d_acc = _dev(torch, pos), _dev(torch, vel), _dev(torch, accel)
    d_img, d_net = torch.zeros((N, 3), dtype=torch.int32, device="cuda"), _dev(torch, net)
    assert Lb.b200_nve_step_one(d_pos.data_ptr(), d_vel.data_ptr(), d_acc.data_ptr(), d_img.data_ptr(), d_grp.data_ptr(),
                                len(members), C.byref(box), C.c_float(dt), 0, C.c_float(0.0), 0, None) == 0
    assert Lb.b200_nve_step_two(d_vel.data_ptr(), d_acc.data_ptr(), d_grp.data_ptr(), len(members), d_net.data_ptr(),
                                C.c_float(dt), 0, C.c_float(0.0), 0, None) == 0
    torch.cuda.synchronize()
    for name, t, orig in (("pos", d_pos, pos), ("vel", d_vel, vel), ("accel", d_acc, accel)):
        got = t.cpu().numpy()
        assert np.array_equal(got.view(np.uint32), p[name].view(np.uint32)), name
        assert np.array_equal(got[~mask].view(np.uint32), orig[~mask].view(np.uint32)), name + ": a non-member moved"
        assert not np.array_equal(got[mask], orig[mask])
    # ---- Langevin step two: per-type gamma, then dscale, both with the tally ----
    for kw in (dict(gamma=[1.0, 2.5]), dict(dscale=0.7)):
        r = reference("langevin", tally=True, **kw)
        r.step_two(123, net)
        p = r.particles()
        d_vel, d_acc = _dev(torch, vel), _dev(torch, accel)
        d_gamma = _dev(torch, np.asarray(kw.get("gamma", [1.0, 1.0]), np.float32))
        d_sum = torch.zeros(1, dtype=torch.float32, device="cuda")
        d_diam, d_tag = _dev(torch, diam), _dev(torch, np.arange(N, dtype=np.uint32))
        la = _lib.LangevinArgs()
        la.d_gamma, la.n_types, la.use_lambda, la.lambda_ = d_gamma.data_ptr(), 2, int("dscale" in kw), kw.get("dscale", 0.0)
        la.T, la.timestep, la.seed, la.d_sum_bdenergy = 1.3, 123, Lb.b200_langevin_hash_seed(11), d_sum.data_ptr()
        la.noiseless_t, la.tally = 0, 1
        assert Lb.b200_langevin_step_two(d_pos.data_ptr(), d_vel.data_ptr(), d_acc.data_ptr(), d_diam.data_ptr(),
                                         d_tag.data_ptr(), d_grp.data_ptr(), len(members), d_net.data_ptr(), C.byref(la),
                                         C.c_float(dt), 3, None) == 0
        torch.cuda.synchronize()
        for name, t, orig in (("vel", d_vel, vel), ("accel", d_acc, accel)):
            got = t.cpu().numpy()
            assert np.array_equal(got.view(np.uint32), p[name].view(np.uint32)), (name, kw)
            assert np.array_equal(got[~mask].view(np.uint32), orig[~mask].view(np.uint32)), name + ": a non-member moved"
        # after one step: reservoir = -E dt + (1/2) E dt, E = sum_j F_bd . v'   =>  E = -2 reservoir / dt
        E_ref = -2.0 * r.langevin_reservoir_energy() / dt
        E = float(d_sum.cpu().numpy()[0])
        assert abs(E - E_ref) <= 5e-4 * abs(E_ref) + 0.05, (E, E_ref, kw)   # FP32 sums of ~1000 terms in different orders


def test_sfc_drivers_vs_reference_sorter(ref, env):
    """b200_sfc_generate_sorted_order + b200_sfc_apply_sorted_order on caller buffers == SFCPackUpdater"""
    hb, _lib, Lb, torch = env
    N, L = 6000, 18.0
    rng = np.random.default_rng(3)
    pos = hb.workloads.random_system(N, L, seed=4, ntypes=2)
    vel = rng.normal(0, 1, (N, 4)).astype(np.float32)
    r = ref.RefSystem(pos, (L, L, L), ntypes=2, vel4=vel)
    r.setup_lj(r_cut=2.0, r_buff=0.3, full=True)
    r.sort_now()
    p = r.particles()
    box = _lib.make_box(L, L, L)
    d_pos, d_vel = _dev(torch, pos), _dev(torch, vel)
    d_order, d_keys = torch.zeros(N, dtype=torch.int32, device="cuda"), torch.zeros(N, dtype=torch.int32, device="cuda")
    d_scr = torch.zeros(2 * N, dtype=torch.int32, device="cuda")
    need = C.c_uint64(0)
    assert Lb.b200_sfc_generate_sorted_order(d_order.data_ptr(), d_keys.data_ptr(), d_scr.data_ptr(), None, 0, C.byref(need),
                                             d_pos.data_ptr(), N, C.byref(box), 256, None) == 0
    d_tmp = torch.zeros(int(need.value) + 16, dtype=torch.uint8, device="cuda")
    assert Lb.b200_sfc_generate_sorted_order(d_order.data_ptr(), d_keys.data_ptr(), d_scr.data_ptr(), d_tmp.data_ptr(),
                                             d_tmp.numel(), None, d_pos.data_ptr(), N, C.byref(box), 256, None) == 0
    torch.cuda.synchronize()
    order = _u32(d_order)
    assert np.array_equal(order, p["tag"])       # new index i holds old particle order[i] == the reference's tag array
    tag = _dev(torch, np.arange(N, dtype=np.uint32))
    o_pos, o_vel, o_tag = torch.zeros_like(d_pos), torch.zeros_like(d_vel), torch.zeros_like(tag)
    d_rtag = torch.zeros(N, dtype=torch.int32, device="cuda")
    ain, aout = _lib.PDataArrays(), _lib.PDataArrays()
    ain.pos, ain.vel, ain.tag = d_pos.data_ptr(), d_vel.data_ptr(), tag.data_ptr()
    aout.pos, aout.vel, aout.tag = o_pos.data_ptr(), o_vel.data_ptr(), o_tag.data_ptr()
    assert Lb.b200_sfc_apply_sorted_order(C.byref(aout), C.byref(ain), d_rtag.data_ptr(), d_order.data_ptr(), N, None) == 0
    torch.cuda.synchronize()
    assert np.array_equal(o_pos.cpu().numpy().view(np.uint32), p["pos"].view(np.uint32))
    assert np.array_equal(o_vel.cpu().numpy().view(np.uint32), p["vel"].view(np.uint32))
    assert np.array_equal(_u32(o_tag), p["tag"]) and np.array_equal(_u32(d_rtag), p["rtag"])
