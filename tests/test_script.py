"""The reference's job-script surface (hoomd_b200/script.py): README lj.py runs unmodified; names, argument meaning and
error behaviour follow hoomd/md/nlist.py, pair.py, integrate.py."""
import numpy as np
import pytest

README_LJ = """
import hoomd
from hoomd import md
hoomd.context.initialize()

# create a 10x10x10 square lattice of particles with name A
hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=2.0, type_name='A'), n=10)
# specify Lennard-Jones interactions between particle pairs
nl = md.nlist.cell()
lj = md.pair.lj(r_cut=3.0, nlist=nl)
lj.pair_coeff.set('A', 'A', epsilon=1.0, sigma=1.0)
# integrate at constant temperature
all = hoomd.group.all();
md.integrate.mode_standard(dt=0.005)
hoomd.md.integrate.langevin(group=all, kT=1.2, seed=4)
# run 10,000 time steps
hoomd.run(NSTEPS)
"""


def _fresh():
    from hoomd_b200 import script
    script.install()
    script.context.initialize()
    return script


def test_surface_and_errors_without_gpu():
    hoomd = _fresh()
    with pytest.raises(RuntimeError):
        hoomd.md.nlist.cell()  # before initialization, like the reference
    sysd = hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=2.0, type_name="A"), n=10)
    assert len(sysd.particles) == 1000 and list(sysd.particles.types) == ["A"]
    snap = sysd.take_snapshot()
    # 10x10x10 sc lattice centred in L = 20: first particle at -9, z fastest (SnapshotParticleData::replicate order)
    assert snap.box.Lx == 20.0
    np.testing.assert_allclose(snap.particles.position[0], [-9, -9, -9])
    np.testing.assert_allclose(snap.particles.position[1], [-9, -9, -7])
    np.testing.assert_allclose(snap.particles.position[-1], [9, 9, 9])
    with pytest.raises(RuntimeError):
        hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=2.0), n=2)  # Cannot initialize more than once
    nl = hoomd.md.nlist.cell(r_buff=0.3, check_period=2)
    lj = hoomd.md.pair.lj(r_cut=2.5, nlist=nl)
    assert nl.r_buff == 0.3 and nl.check_period == 2 and lj in nl.subscribers
    with pytest.raises(RuntimeError):
        lj.set_params(mode="bogus")
    lj.set_params(mode="shift")
    assert lj.shift_mode == 1
    # missing coefficients are an error at run time (update_coeffs), not at construction
    with pytest.raises(RuntimeError):
        lj.tables(["A"])
    lj.pair_coeff.set("A", "A", epsilon=1.5, sigma=0.9)
    t = lj.tables(["A"])
    assert t["epsilon"][0, 0] == 1.5 and t["alpha"][0, 0] == 1.0 and t["r_cut"][0, 0] == 2.5
    lj.pair_coeff.set("A", "A", r_cut=2.0)
    assert lj.tables(["A"])["r_cut"][0, 0] == 2.0
    hoomd.md.integrate.mode_standard(dt=0.005)
    lv = hoomd.md.integrate.langevin(group=hoomd.group.all(), kT=1.2, seed=4)
    lv.set_gamma("A", 2.0)
    with pytest.raises(RuntimeError):
        lv.set_gamma("B", 1.0)


def test_two_type_coefficient_table():
    hoomd = _fresh()
    snap = hoomd.data.make_snapshot(N=4, box=hoomd.data.boxdim(L=10), particle_types=["A", "B"])
    snap.particles.typeid[:] = [0, 1, 0, 1]
    hoomd.init.read_snapshot(snap)
    nl = hoomd.md.nlist.cell()
    lj = hoomd.md.pair.lj(r_cut=2.5, nlist=nl)
    lj.pair_coeff.set("A", "A", epsilon=1.0, sigma=1.0)
    lj.pair_coeff.set("A", "B", epsilon=1.5, sigma=0.8, r_cut=2.0)
    with pytest.raises(RuntimeError):
        lj.tables(["A", "B"])  # (B, B) missing
    lj.pair_coeff.set("B", "B", epsilon=0.5, sigma=0.88, r_cut=2.2)
    t = lj.tables(["A", "B"])
    assert t["epsilon"][0, 1] == t["epsilon"][1, 0] == 1.5
    assert t["r_cut"][1, 1] == 2.2 and t["r_cut"][0, 1] == 2.0 and t["r_cut"][0, 0] == 2.5


@pytest.mark.gpu
def test_readme_lj_runs_unmodified_and_matches_reference():
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref/libhoomd_ref.so not built")
    hoomd = _fresh()
    nsteps = 60
    ns = {}
    exec(README_LJ.replace("NSTEPS", str(nsteps)), ns)
    snap = hoomd.context.current.system.take_snapshot()
    assert hoomd.get_step() == nsteps
    # the same script on the reference's CPU classes
    from hoomd_b200 import workloads
    pos, vel, L = workloads.sc_lattice(10, 2.0)
    assert L == 20.0
    r = R.RefSystem(pos, (L, L, L), vel4=vel)
    r.setup_lj(r_cut=3.0, r_buff=0.4, full=True)
    r.setup_integrator(0.005)
    r.add_langevin(1.2, 4)
    r.set_sorter(300)
    r.run(nsteps)
    a = r.particles()
    order = np.argsort(a["tag"])
    # identical lattice order => tags coincide
    assert np.abs(a["pos"][order][:, :3] - snap.particles.position).max() < 2e-4
    assert np.abs(a["vel"][order][:, :3] - snap.particles.velocity).max() < 2e-3
    # logging surface
    log = hoomd.analyze.log(filename=None, quantities=["temperature", "potential_energy"], period=10)
    hoomd.run(20)
    assert len(log.rows) == 2 and 0.0 < log.rows[0][1] < 5.0


KG_SCRIPT = """
import hoomd
from hoomd import md
hoomd.context.initialize()
snap = hoomd.data.make_snapshot(N=N, box=hoomd.data.boxdim(L=L), particle_types=['A'], bond_types=['polymer'])
snap.particles.position[:] = POS
snap.particles.velocity[:] = VEL
snap.bonds.resize(len(BONDS))
snap.bonds.group[:] = BONDS
system = hoomd.init.read_snapshot(snap)
nl = md.nlist.cell()
wca = md.pair.lj(r_cut=2**(1.0/6.0), nlist=nl)
wca.pair_coeff.set('A', 'A', epsilon=1.0, sigma=1.0)
wca.set_params(mode='shift')
fene = md.bond.fene()
fene.bond_coeff.set('polymer', k=30.0, r0=1.5, sigma=1.0, epsilon=1.0)
md.integrate.mode_standard(dt=0.005)
md.integrate.nve(group=hoomd.group.all())
hoomd.run(NSTEPS)
"""


def test_bond_surface_without_gpu():
    hoomd = _fresh()
    snap = hoomd.data.make_snapshot(N=4, box=hoomd.data.boxdim(L=10), particle_types=["A"], bond_types=["polymer"])
    hoomd.init.read_snapshot(snap)
    with pytest.raises(RuntimeError):
        hoomd.md.bond.fene()  # "No bonds are defined." (hoomd/md/bond.py:348-351)
    sysd = hoomd.context.current.system
    assert sysd.bonds.add("polymer", 0, 1) == 0 and len(sysd.bonds) == 1
    with pytest.raises(RuntimeError):
        sysd.bonds.add("bogus", 1, 2)
    fene = hoomd.md.bond.fene()
    with pytest.raises(RuntimeError):
        fene.bond_coeff.table(["polymer"], fene.required_coeffs)  # coefficients not set
    fene.bond_coeff.set("polymer", k=30.0, r0=1.5, sigma=1.0, epsilon=1.0)
    assert fene.bond_coeff.table(["polymer"], fene.required_coeffs)["k"] == [30.0]
    nl = hoomd.md.nlist.cell()
    assert nl.exclusions == ["bond", "body"]
    nl.reset_exclusions([])
    assert nl.exclusions == []
    with pytest.raises(RuntimeError):
        nl.reset_exclusions(["1-4"])
    s2 = sysd.take_snapshot()
    assert s2.bonds.N == 1 and list(s2.bonds.group[0]) == [0, 1] and s2.bonds.types == ["polymer"]


@pytest.mark.gpu
def test_kremer_grest_script_matches_reference():
    """a Kremer-Grest job script (snapshot with bonds, WCA via pair.lj shift, bond.fene, default bond exclusions, NVE)"""
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref/libhoomd_ref.so not built")
    from hoomd_b200 import workloads
    hoomd = _fresh()
    pos, vel, L, bonds = workloads.polymer_lattice(8, 16, rho=0.85, kT=1.0, seed=21)
    nsteps = 80
    ns = dict(N=len(pos), L=L, POS=pos[:, :3], VEL=vel[:, :3], BONDS=bonds, NSTEPS=nsteps)
    exec(KG_SCRIPT, ns)
    snap = hoomd.context.current.system.take_snapshot()
    r = R.RefSystem(pos, (L, L, L), vel4=vel, n_bond_types=1)
    r.set_bonds(bonds)
    r.setup_lj(r_cut=2.0 ** (1.0 / 6.0), r_buff=0.4, full=True, shift_mode=1)
    r.setup_fene(k=30.0, r0=1.5)
    r.exclusions_from_bonds(True)
    r.setup_integrator(0.005)
    r.add_nve()
    r.set_sorter(300)
    r.run(nsteps)
    a = r.particles()
    order = np.argsort(a["tag"])
    assert np.abs(a["pos"][order][:, :3] - snap.particles.position).max() < 2e-4
    assert np.abs(a["vel"][order][:, :3] - snap.particles.velocity).max() < 2e-3
    assert snap.bonds.N == len(bonds)
    assert abs(ns["fene"].get_energy() - 20.0 * len(bonds)) < 5.0 * len(bonds)


NVT_SCRIPT = """
import hoomd
from hoomd import md
hoomd.context.initialize()
hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=1.6, type_name='A'), n=9)
nl = md.nlist.cell()
lj = md.pair.lj(r_cut=2.5, nlist=nl)
lj.pair_coeff.set('A', 'A', epsilon=1.0, sigma=1.0)
md.integrate.mode_standard(dt=0.005)
nvt = md.integrate.nvt(group=hoomd.group.all(), kT=1.0, tau=0.5)
hoomd.run(NSTEPS)
"""


def test_nvt_surface_without_gpu():
    hoomd = _fresh()
    from hoomd import md
    hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=2.0, type_name="A"), n=4)
    m = md.integrate.nvt(group=hoomd.group.all(), kT=1.2, tau=0.5)
    m.set_params(kT=1.5)
    m.set_params(tau=1.0)
    assert (m.kT, m.tau) == (1.5, 1.0)
    with pytest.raises(TypeError):
        md.integrate.nvt(group=hoomd.group.all(), kT=1.0)  # tau is required (hoomd/md/integrate.py:201)


@pytest.mark.gpu
def test_nvt_script_matches_reference():
    """md.integrate.nvt from a job script: 80 steps against TwoStepNVTMTK on the reference's CPU classes; a second
    hoomd.run continues with the thermostat state, also across an engine re-configuration (set_params)"""
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref/libhoomd_ref.so not built")
    hoomd = _fresh()
    ns = {}
    exec(NVT_SCRIPT.replace("NSTEPS", "80"), ns)
    snap = hoomd.context.current.system.take_snapshot()
    from hoomd_b200 import workloads
    pos, vel, L = workloads.sc_lattice(9, 1.6)
    r = R.RefSystem(pos, (L, L, L), vel4=vel)
    r.setup_lj(r_cut=2.5, r_buff=0.4, full=True)
    r.setup_integrator(0.005)
    r.add_nvt(1.0, 0.5)
    r.set_sorter(300)
    r.run(80)
    a = r.particles()
    order = np.argsort(a["tag"])
    assert np.abs(a["pos"][order][:, :3] - snap.particles.position).max() < 2e-4
    assert np.abs(a["vel"][order][:, :3] - snap.particles.velocity).max() < 2e-3
    eng = hoomd.context.current.engine
    assert np.allclose(eng.nvt_state(), r.nvt_state(), rtol=2e-5, atol=1e-7)
    log = hoomd.analyze.log(filename=None, quantities=["nvt_mtk_reservoir_energy_all", "temperature"], period=10)
    assert abs(log.query("nvt_mtk_reservoir_energy_all") - r.nvt_reservoir_energy()) <= 1e-4 * abs(r.nvt_reservoir_energy())
    hoomd.context.current.loggers.remove(log)
    hoomd.context.current.engine_dirty = False
    # continue; the re-configured engine must pick the thermostat variables up again
    ns["nvt"].set_params(tau=0.5)
    hoomd.run(20)
    r.run(20)
    eng = hoomd.context.current.engine
    assert np.allclose(eng.nvt_state(), r.nvt_state(), rtol=1e-4, atol=1e-6), (eng.nvt_state(), r.nvt_state())


BOX_SCRIPT = """
import hoomd
from hoomd import md
hoomd.context.initialize()
hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=1.8, type_name='A'), n=9)
nl = md.nlist.cell()
lj = md.pair.lj(r_cut=2.5, nlist=nl)
lj.pair_coeff.set('A', 'A', epsilon=1.0, sigma=1.0)
md.integrate.mode_standard(dt=0.005)
md.integrate.langevin(group=hoomd.group.all(), kT=1.0, seed=5)
resize = hoomd.update.box_resize(L=hoomd.variant.linear_interp([(0, L0), (100, L1)]), period=10)
hoomd.run(120)
"""


def test_variant_and_box_resize_surface_without_gpu():
    hoomd = _fresh()
    hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=2.0, type_name="A"), n=4)
    v = hoomd.variant.linear_interp([(0, 10.0), (100, 20.0)])
    assert v.getValue(0) == 10.0 and v.getValue(50) == 15.0 and v.getValue(100) == 20.0 and v.getValue(10 ** 6) == 20.0
    assert hoomd.variant.linear_interp([(10, 1.0), (20, 3.0)]).getValue(5) == 1.0
    with pytest.raises(RuntimeError):
        hoomd.variant.linear_interp([])
    with pytest.raises(RuntimeError):
        hoomd.variant.linear_interp([(0, 1.0)], zero=5)  # zero in the future
    with pytest.raises(RuntimeError):
        hoomd.update.box_resize()  # nothing to resize (hoomd/update.py:344-346)
    with pytest.raises(RuntimeError):
        hoomd.update.box_resize(xy=0.5)  # triclinic: out of scope
    # scheduling as System::addUpdater / updater_item::shouldExecute (hoomd/System.cc:212-243, System.h:205-242)
    br = hoomd.update.box_resize(L=hoomd.variant.linear_interp([(0, 8.0), (100, 6.0)]), period=10)
    assert br.next_execute == 0 and br.should_execute(0) and br.next_execute == 10
    assert not br.should_execute(5) and br.steps_until(3) == 7 and br.steps_until(10) is None
    # phase: start_step = ceil(cur / period) * period + phase, then += period: 3, 13, 23 (not 7, 17)
    b3 = hoomd.update.box_resize(L=7.0, period=10, phase=3)
    fired = [t for t in range(30) if b3.should_execute(t)]
    assert fired == [3, 13, 23]
    lg = hoomd.analyze.log(filename=None, quantities=["temperature"], period=5, phase=2)
    assert [t for t in range(20) if lg.should_execute(t)] == [2, 7, 12, 17]
    b3.disable(); lg.disable()
    # phase -1: starts on the current step
    bm = hoomd.update.box_resize(L=7.0, period=4, phase=-1)
    assert [t for t in range(10) if bm.should_execute(t)] == [0, 4, 8]


@pytest.mark.gpu
def test_box_resize_script_matches_reference():
    """update.box_resize(L=variant.linear_interp(...), period=10) from a job script against BoxResizeUpdater driven the
    same way on the reference's CPU classes (sorter period 300 on both)"""
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref/libhoomd_ref.so not built")
    from hoomd_b200 import workloads
    pos, vel, L = workloads.sc_lattice(9, 1.8)
    L1 = 0.9 * L
    hoomd = _fresh()
    ns = {}
    exec(BOX_SCRIPT.replace("L0", repr(float(L))).replace("L1", repr(float(L1))), ns)
    snap = hoomd.context.current.system.take_snapshot()
    r = R.RefSystem(pos, (L, L, L), vel4=vel)
    r.setup_lj(r_cut=2.5, r_buff=0.4, full=True)
    r.setup_integrator(0.005)
    r.add_langevin(1.0, 5)
    r.set_sorter(300)
    for ts in range(0, 120, 10):
        f = min(ts, 100) / 100.0
        Lt = np.float32((1.0 - f) * float(L) + f * float(L1)) if ts < 100 else np.float32(L1)
        r.set_box(Lt, Lt, Lt, True)
        r.run(10)
    assert abs(hoomd.context.current.system.box.Lx - float(np.float32(L1))) < 1e-6
    assert np.array_equal(r.box_lengths(), hoomd.context.current.engine.box_lengths())
    a = r.particles()
    order = np.argsort(a["tag"])
    d = a["pos"][order][:, :3] - snap.particles.position
    d -= np.float32(L1) * np.round(d / np.float32(L1))
    assert np.abs(d).max() < 3e-4
    assert np.abs(a["vel"][order][:, :3] - snap.particles.velocity).max() < 3e-3


def test_nlist_dist_check_is_forwarded():
    hoomd = _fresh()
    from hoomd import md
    hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=2.0, type_name="A"), n=4)
    nl = md.nlist.cell(check_period=5, dist_check=False)
    assert nl.dist_check is False and nl.check_period == 5
    nl.set_params(dist_check=True)
    assert nl.dist_check is True
    import inspect
    from hoomd_b200 import script
    assert "set_dist_check(lj.nlist.dist_check)" in inspect.getsource(script._configure_engine)


def test_dpd_conservative_surface_without_gpu():
    hoomd = _fresh()
    from hoomd import md
    hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=1.0, type_name="A"), n=4)
    nl = md.nlist.cell()
    dpdc = md.pair.dpd_conservative(r_cut=1.0, nlist=nl)
    dpdc.pair_coeff.set("A", "A", A=25.0)
    t = dpdc.tables(["A"])
    assert t["A"][0][0] == 25.0 and t["r_cut"][0][0] == 1.0
    with pytest.raises(RuntimeError):
        dpdc.set_params(mode="shift")  # hoomd/md/pair.py:1502-1511
    from hoomd_b200 import _lib
    p = _lib.pair_params("dpd_conservative", 1, A=25.0)
    assert p.shape == (1, 1, 4) and p[0, 0, 0] == 25.0 and p[0, 0, 1] == 0.0  # (A, gamma = 0)


def test_temperature_variant_surface_without_gpu():
    hoomd = _fresh()
    from hoomd import md
    hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=2.0, type_name="A"), n=4)
    ramp = hoomd.variant.linear_interp([(0, 2.0), (100, 1.0)])
    lv = md.integrate.langevin(group=hoomd.group.all(), kT=ramp, seed=1)
    assert lv.kT == 2.0 and lv.kT_is_variable
    lv.set_params(kT=1.5)
    assert lv.kT == 1.5 and not lv.kT_is_variable
    lv.disable()
    nvt = md.integrate.nvt(group=hoomd.group.all(), kT=ramp, tau=0.5)
    assert nvt.kT_is_variable and nvt.kT == 2.0


@pytest.mark.gpu
def test_temperature_ramp_script_matches_reference():
    """md.integrate.langevin(kT=variant.linear_interp(...)): the set point follows the variant step by step"""
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref/libhoomd_ref.so not built")
    from hoomd_b200 import workloads
    hoomd = _fresh()
    from hoomd import md
    hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=1.6, type_name="A"), n=9)
    nl = md.nlist.cell()
    lj = md.pair.lj(r_cut=2.5, nlist=nl)
    lj.pair_coeff.set("A", "A", epsilon=1.0, sigma=1.0)
    md.integrate.mode_standard(dt=0.005)
    md.integrate.langevin(group=hoomd.group.all(), kT=hoomd.variant.linear_interp([(0, 2.0), (50, 0.8)]), seed=3)
    hoomd.run(60)
    snap = hoomd.context.current.system.take_snapshot()
    pos, vel, L = workloads.sc_lattice(9, 1.6)
    r = R.RefSystem(pos, (L, L, L), vel4=vel)
    r.setup_lj(r_cut=2.5, r_buff=0.4, full=True)
    r.setup_integrator(0.005)
    r.add_langevin(2.0, 3)
    r.set_sorter(300)
    for ts in range(60):
        f = min(ts, 50) / 50.0
        r.set_kT(np.float32((1.0 - f) * 2.0 + f * 0.8))
        r.run(1)
    a = r.particles()
    order = np.argsort(a["tag"])
    assert np.abs(a["pos"][order][:, :3] - snap.particles.position).max() < 2e-4
    assert np.abs(a["vel"][order][:, :3] - snap.particles.velocity).max() < 2e-3


def test_log_file_format(tmp_path):
    """analyze.log writes the reference's plain-text format (hoomd/LogPlainTXT.cc): header for a new file, tab-delimited
    rows with 10 significant digits, appending to an existing file unless overwrite=True"""
    hoomd = _fresh()
    hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=2.0, type_name="A"), n=4)
    fn = str(tmp_path / "log.txt")
    lg = hoomd.analyze.log(filename=fn, quantities=["temperature", "potential_energy"], period=10, header_prefix="#")
    lg._write([0, 1.25, -3.0])
    lg._write([10, 1.0 / 3.0, 2.5e-7])
    lines = open(fn).read().splitlines()
    assert lines[0] == "#timestep\ttemperature\tpotential_energy"
    assert lines[1] == "0\t1.25\t-3" and lines[2] == "10\t0.3333333333\t2.5e-07"
    lg2 = hoomd.analyze.log(filename=fn, quantities=["temperature"], period=10)
    lg2._write([20, 2.0])
    assert open(fn).read().splitlines()[-1] == "20\t2" and len(open(fn).read().splitlines()) == 4
    lg3 = hoomd.analyze.log(filename=fn, quantities=["temperature"], period=10, overwrite=True)
    lg3._write([0, 1.0])
    assert open(fn).read().splitlines() == ["timestep\ttemperature", "0\t1"]


@pytest.mark.gpu
def test_first_logged_row_sees_real_forces_and_per_force_energies():
    """The reference computes forces in prepRun before the step loop, so analyze.log's row at the first step holds the real
    potential energy and a pressure with its virial term; pair_lj_energy / bond_fene_energy are the energies of ONE
    ForceCompute each (ForceCompute::calcEnergySum), not the total."""
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref/libhoomd_ref.so not built")
    from hoomd_b200 import workloads
    hoomd = _fresh()
    from hoomd import md
    pos, vel, L, bonds = workloads.polymer_lattice(6, 8, rho=0.85, kT=1.0, seed=3)
    pos[:, :3] += np.random.default_rng(1).normal(0, 0.03, (len(pos), 3)).astype(np.float32)
    snap = hoomd.data.make_snapshot(N=len(pos), box=hoomd.data.boxdim(L=L), particle_types=["A"], bond_types=["polymer"])
    snap.particles.position[:] = pos[:, :3]
    snap.particles.velocity[:] = vel[:, :3]
    snap.bonds.resize(len(bonds))
    snap.bonds.group[:] = bonds
    hoomd.init.read_snapshot(snap)
    nl = md.nlist.cell()
    wca = md.pair.lj(r_cut=2.0 ** (1.0 / 6.0), nlist=nl)
    wca.pair_coeff.set("A", "A", epsilon=1.0, sigma=1.0)
    wca.set_params(mode="shift")
    fene = md.bond.fene()
    fene.bond_coeff.set("polymer", k=30.0, r0=1.5, sigma=1.0, epsilon=1.0)
    md.integrate.mode_standard(dt=0.005)
    md.integrate.nve(group=hoomd.group.all())
    log = hoomd.analyze.log(filename=None, period=10, quantities=["potential_energy", "pressure", "pair_lj_energy",
                                                                 "bond_fene_energy", "kinetic_energy"])
    with pytest.raises(RuntimeError):
        hoomd.analyze.log(filename=None, quantities=["bogus_quantity"], period=10)
    hoomd.run(10)
    row0 = dict(zip(["ts"] + log.quantities, log.rows[0]))
    assert row0["ts"] == 0 and len(log.rows) == 1
    # the same state on the reference's CPU classes
    r = R.RefSystem(pos, (L, L, L), vel4=vel, n_bond_types=1)
    r.set_bonds(bonds)
    r.setup_lj(r_cut=2.0 ** (1.0 / 6.0), r_buff=0.4, full=True, shift_mode=1)
    r.setup_fene(k=30.0, r0=1.5)
    r.exclusions_from_bonds(True)
    r.set_virial(True)
    r.setup_integrator(0.005)
    r.add_nve()
    r.run(0)  # prepRun only
    th = r.thermo()
    pe_pair = float(r.force_component(0)[0][:, 3].astype(np.float64).sum())
    pe_bond = float(r.force_component(1)[0][:, 3].astype(np.float64).sum())
    assert pe_bond > 1.0 and abs(pe_pair) > 1e-3  # both contribute in this state
    assert abs(row0["potential_energy"] - th["pe"]) <= 2e-5 * abs(th["pe"])
    assert abs(row0["pressure"] - th["P"]) <= 2e-5 * abs(th["P"])
    assert abs(row0["kinetic_energy"] - th["ke"]) <= 2e-5 * abs(th["ke"])
    assert abs(row0["pair_lj_energy"] - pe_pair) <= 2e-5 * abs(pe_pair) + 1e-6
    assert abs(row0["bond_fene_energy"] - pe_bond) <= 2e-5 * abs(pe_bond)
    assert abs(row0["pair_lj_energy"] + row0["bond_fene_energy"] - row0["potential_energy"]) <= 1e-4 * abs(th["pe"])
    assert abs(wca.get_energy() + fene.get_energy() - hoomd.context.current.engine.thermo()["pe"]) \
        <= 1e-4 * abs(th["pe"])
    with pytest.raises(RuntimeError):
        log.query("pair_gauss_energy")  # no such force


def test_gsd_surface_without_gpu(tmp_path):
    """dump.gsd(period=None) writes the initial configuration without an engine; init.read_gsd / data.gsd_snapshot read it back,
    restart= takes precedence when that file exists, the time step comes from the file"""
    from hoomd_b200 import gsd_io
    hoomd = _fresh()
    sysd = hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=2.0, type_name="A"), n=4)
    snap = sysd.take_snapshot()
    snap.particles.velocity[:] = np.random.default_rng(1).normal(0, 1, (64, 3))
    snap.particles.mass[:] = 2.0
    sysd.restore_snapshot(snap)
    init_file, restart_file = str(tmp_path / "init.gsd"), str(tmp_path / "restart.gsd")
    with pytest.raises(ValueError):
        hoomd.dump.gsd(init_file, period=None, group=hoomd.group.all(), static=["attribute"], dynamic=["momentum"])
    hoomd.dump.gsd(init_file, period=None, group=hoomd.group.all(), overwrite=True, time_step=77)
    fr = gsd_io.read_hoomd_frame(init_file)
    assert fr["step"] == 77 and fr["N"] == 64 and np.allclose(fr["box"][:3], 8.0) and (fr["mass"] == 2.0).all()
    assert np.array_equal(fr["velocity"], snap.particles.velocity)
    s2 = hoomd.data.gsd_snapshot(init_file)
    assert np.array_equal(s2.particles.position, snap.particles.position) and s2.box.Lx == 8.0
    # a sub-group: N of the file is the group size, no topology
    hoomd.dump.gsd(str(tmp_path / "grp.gsd"), period=None, group=hoomd.group.tags(10, 19), overwrite=True)
    g = gsd_io.read_hoomd_frame(str(tmp_path / "grp.gsd"))
    assert g["N"] == 10 and np.array_equal(g["position"], snap.particles.position[10:20])
    # read_gsd: restart file wins when present
    hoomd = _fresh()
    hoomd.init.read_gsd(init_file, restart=restart_file)
    assert hoomd.context.current.timestep == 77 and hoomd.get_step() == 77
    hoomd = _fresh()
    w = gsd_io.HoomdGSDWriter(restart_file, overwrite=True, truncate=True)
    w.write_frame(500, (9, 9, 9), dict(position=np.zeros((5, 3), np.float32)))
    w.close()
    sysd = hoomd.init.read_gsd(init_file, restart=restart_file)
    assert hoomd.context.current.timestep == 500 and len(sysd.particles) == 5 and sysd.box.Lx == 9.0
    hoomd = _fresh()
    hoomd.init.read_gsd(init_file, time_step=3)
    assert hoomd.context.current.timestep == 3


@pytest.mark.gpu
def test_gsd_trajectory_and_restart(tmp_path):
    """dump.gsd as an analyzer on the device-resident state: frames at the scheduled steps, and a job restarted from its
    truncate=True restart file continues the trajectory (NVE: a(t) = F(t)/m is recomputed from the restored positions, as the
    reference does for a file without accelerations)."""
    from hoomd_b200 import gsd_io
    traj, restart = str(tmp_path / "traj.gsd"), str(tmp_path / "restart.gsd")

    def setup(hoomd):
        nl = hoomd.md.nlist.cell()
        lj = hoomd.md.pair.lj(r_cut=2.5, nlist=nl)
        lj.pair_coeff.set("A", "A", epsilon=1.0, sigma=1.0)
        hoomd.md.integrate.mode_standard(dt=0.005)
        hoomd.md.integrate.nve(group=hoomd.group.all())

    hoomd = _fresh()
    hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=1.5, type_name="A"), n=8)
    snap = hoomd.context.current.system.take_snapshot()
    snap.particles.velocity[:] = np.random.default_rng(2).normal(0, 1.0, (512, 3))
    hoomd.context.current.system.restore_snapshot(snap)
    setup(hoomd)
    hoomd.dump.gsd(traj, period=40, group=hoomd.group.all(), overwrite=True, dynamic=["momentum"])
    rs = hoomd.dump.gsd(restart, period=100, group=hoomd.group.all(), overwrite=True, truncate=True, phase=0)
    hoomd.run(100)
    rs.write_restart()                       # frame at step 100
    mid = hoomd.context.current.system.take_snapshot()
    hoomd.run(60)
    end = hoomd.context.current.system.take_snapshot()
    with gsd_io.GSDFile(traj) as f:
        assert f.nframes == 4                # steps 0, 40, 80, 120 (160 is the first step of the NEXT run)
        assert [int(f.read_chunk(k, "configuration/step")[0, 0]) for k in range(4)] == [0, 40, 80, 120]
        assert f.find_chunk(2, "particles/velocity") is not None and f.find_chunk(2, "particles/mass") is None
    fr = gsd_io.read_hoomd_frame(restart)
    assert fr["step"] == 100
    assert np.array_equal(fr["position"], mid.particles.position) and np.array_equal(fr["image"], mid.particles.image)

    # second job: restart from the file
    hoomd = _fresh()
    hoomd.init.read_gsd(traj, restart=restart)
    assert hoomd.get_step() == 100
    setup(hoomd)
    hoomd.run(60)
    again = hoomd.context.current.system.take_snapshot()
    assert hoomd.get_step() == 160
    # same state to FP32 round-off growth: the restarted job holds its particles in tag order until the next sort, so the
    # neighbor rows (and with them the order of the force sums) differ from the uninterrupted job's
    assert np.abs(again.particles.position - end.particles.position).max() < 2e-5
    assert np.abs(again.particles.velocity - end.particles.velocity).max() < 2e-4
    assert np.array_equal(again.particles.image, end.particles.image)
