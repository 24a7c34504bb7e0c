"""The reference's Python job-script surface for the MD hot path, driving the B200 engine.

An unmodified HOOMD-blue 2.3 job script of the shape in the reference README (README.md:56-72)

    import hoomd
    from hoomd import md
    hoomd.context.initialize()
    hoomd.init.create_lattice(unitcell=hoomd.lattice.sc(a=2.0, type_name='A'), n=10)
    nl = md.nlist.cell()
    lj = md.pair.lj(r_cut=3.0, nlist=nl)
    lj.pair_coeff.set('A', 'A', epsilon=1.0, sigma=1.0)
    all = hoomd.group.all()
    md.integrate.mode_standard(dt=0.005)
    md.integrate.langevin(group=all, kT=1.2, seed=4)
    hoomd.run(10e3)

runs after ``import hoomd_b200.script as hoomd`` (or ``hoomd_b200.script.install()``, which registers the modules as
``hoomd`` / ``hoomd.md``).  Same names, argument meaning and error behaviour as hoomd/md/nlist.py:555,
hoomd/md/pair.py:358-445,579-609, hoomd/md/integrate.py:49,729,844, hoomd/__init__.py (run), hoomd/init.py:32-93,
hoomd/data.py (make_snapshot / take_snapshot, snapshot.bonds, system.bonds.add), hoomd/md/bond.py:236-372 (bond.fene,
bond.harmonic), hoomd/md/nlist.py:213-285 (reset_exclusions), hoomd/update.py:239-263 (sort).  Like the reference, commands only record
parameters; the C++/CUDA objects are (re)configured when ``run()`` starts (hoomd/md/nlist.py:80-109 update_rcut,
hoomd/md/pair.py:430-445 update_coeffs).  Only the path of SURVEY.md section 8 is covered: one pair force (lj), one
integration method on group.all(), orthorhombic boxes.
"""
import sys
import types

import numpy as np


class _Context:
    def __init__(self):
        self.reset()

    def reset(self):
        self.system = None          # _SystemData
        self.neighbor_lists = []
        self.forces = []
        self.integrator = None
        self.integration_methods = []
        self.sorter = update.sort() if "update" in globals() else None
        self.loggers = []
        self.dumpers = []
        self.box_resizers = []
        self.engine = None
        self.engine_dirty = True
        self.timestep = 0


class _ContextModule(types.ModuleType):
    """hoomd.context: initialize(), current"""

    def __init__(self):
        super().__init__("hoomd.context")
        self.current = None
        self.exec_conf = None

    def initialize(self, args=None):
        self.current = _Context()
        self.current.sorter = update.sort()
        return self.current


# ---- hoomd.lattice / hoomd.init / hoomd.data -------------------------------------------------------------------------
class _UnitCell:
    def __init__(self, N, a1, a2, a3, position, type_name):
        self.N, self.a1, self.a2, self.a3 = N, np.asarray(a1, float), np.asarray(a2, float), np.asarray(a3, float)
        self.position, self.type_name = np.asarray(position, float), type_name


lattice = types.ModuleType("hoomd.lattice")
lattice.sc = lambda a, type_name="A": _UnitCell(1, [a, 0, 0], [0, a, 0], [0, 0, a], [[0, 0, 0]], [type_name])
lattice.bcc = lambda a, type_name="A": _UnitCell(2, [a, 0, 0], [0, a, 0], [0, 0, a],
                                                 [[0, 0, 0], [a / 2, a / 2, a / 2]], [type_name] * 2)
lattice.fcc = lambda a, type_name="A": _UnitCell(4, [a, 0, 0], [0, a, 0], [0, 0, a],
                                                 [[0, 0, 0], [0, a / 2, a / 2], [a / 2, 0, a / 2], [a / 2, a / 2, 0]],
                                                 [type_name] * 4)


class _BoxDim:
    """hoomd.data.boxdim (orthorhombic)"""

    def __init__(self, Lx=1.0, Ly=1.0, Lz=1.0, L=None, dimensions=3):
        if L is not None:
            Lx = Ly = Lz = L
        self.Lx, self.Ly, self.Lz, self.xy, self.xz, self.yz, self.dimensions = Lx, Ly, Lz, 0.0, 0.0, 0.0, dimensions

    def get_volume(self):
        return self.Lx * self.Ly * self.Lz


class _ParticleSnapshot:
    def __init__(self, N, types):
        self.N = N
        self.types = list(types)
        self.position = np.zeros((N, 3), np.float32)
        self.velocity = np.zeros((N, 3), np.float32)
        self.acceleration = np.zeros((N, 3), np.float32)
        self.typeid = np.zeros(N, np.uint32)
        self.mass = np.ones(N, np.float32)
        self.diameter = np.ones(N, np.float32)
        self.image = np.zeros((N, 3), np.int32)


class _BondSnapshot:
    """snapshot.bonds (hoomd/data.py: N, types, typeid, group, resize)"""

    def __init__(self, types=()):
        self.types = list(types)
        self.resize(0)

    def resize(self, n):
        old_t, old_g = getattr(self, "typeid", np.zeros(0, np.uint32)), getattr(self, "group", np.zeros((0, 2), np.uint32))
        self.N = int(n)
        self.typeid = np.zeros(self.N, np.uint32)
        self.group = np.zeros((self.N, 2), np.uint32)
        k = min(len(old_t), self.N)
        self.typeid[:k], self.group[:k] = old_t[:k], old_g[:k]


class _Snapshot:
    def __init__(self, N, box, particle_types, bond_types=()):
        self.box = box
        self.particles = _ParticleSnapshot(N, particle_types)
        self.bonds = _BondSnapshot(bond_types)


class _ParticleTypes:
    def __init__(self, names):
        self._names = list(names)

    def __len__(self):
        return len(self._names)

    def __getitem__(self, i):
        return self._names[i]

    def __iter__(self):
        return iter(self._names)


class _Bonds:
    """system.bonds (hoomd/data.py bond_data): len(), add(type, a, b)"""

    def __init__(self, sysdata):
        self._s = sysdata

    def __len__(self):
        return len(self._s.bond_group)

    def add(self, type, a, b):
        s = self._s
        if type not in s.bond_types:
            raise RuntimeError("Error adding bond: unknown bond type " + str(type))
        if max(int(a), int(b)) >= s.N:
            raise RuntimeError("Particle tag out of bounds when attempting to add bond")
        s.bond_group = np.concatenate([s.bond_group, np.array([[a, b]], np.uint32)])
        s.bond_typeid = np.concatenate([s.bond_typeid, np.array([s.bond_types.index(type)], np.uint32)])
        context.current.engine_dirty = True
        return len(s.bond_group) - 1


class _Particles:
    def __init__(self, sysdata):
        self._s = sysdata
        self.types = _ParticleTypes(sysdata.types)

    def __len__(self):
        return self._s.N


class _SystemData:
    """hoomd.data.system_data: the host-side snapshot the engine is seeded from / synchronised back into"""

    def __init__(self, snapshot):
        p = snapshot.particles
        self.N = int(p.N)
        self.types = list(p.types)
        self.box = snapshot.box
        self.pos4 = np.zeros((self.N, 4), np.float32)
        self.pos4[:, :3] = p.position
        self.pos4[:, 3] = np.asarray(p.typeid, np.int32).view(np.float32)
        self.vel4 = np.zeros((self.N, 4), np.float32)
        self.vel4[:, :3] = p.velocity
        self.vel4[:, 3] = p.mass
        self.image = np.array(p.image, np.int32)
        self.diameter = np.array(getattr(p, "diameter", np.ones(self.N)), np.float32)
        self.particles = _Particles(self)
        b = getattr(snapshot, "bonds", None) or _BondSnapshot()
        self.bond_types = list(b.types)
        self.bond_group = np.array(b.group, np.uint32).reshape(-1, 2)
        self.bond_typeid = np.array(b.typeid, np.uint32)
        self.bonds = _Bonds(self)

    def take_snapshot(self, **kwargs):
        """current state in TAG order (hoomd/data.py take_snapshot)"""
        ctx = context.current
        if ctx.engine is not None:
            st = ctx.engine.particles()
            order = np.argsort(st["tag"], kind="stable")
            self.pos4, self.vel4, self.image = st["pos"][order], st["vel"][order], st["image"][order]
            accel = st["accel"][order]
        else:
            accel = np.zeros((self.N, 3), np.float32)
        snap = _Snapshot(self.N, self.box, self.types, self.bond_types)
        snap.bonds.resize(len(self.bond_group))
        snap.bonds.group[:] = self.bond_group
        snap.bonds.typeid[:] = self.bond_typeid
        q = snap.particles
        q.position[:] = self.pos4[:, :3]
        q.velocity[:] = self.vel4[:, :3]
        q.mass[:] = self.vel4[:, 3]
        q.typeid[:] = self.pos4[:, 3].view(np.int32)
        q.image[:] = self.image
        q.diameter[:] = self.diameter
        q.acceleration[:] = accel
        return snap

    def restore_snapshot(self, snapshot):
        self.__init__(snapshot)
        context.current.engine_dirty = True


data = types.ModuleType("hoomd.data")
data.boxdim = _BoxDim
data.make_snapshot = lambda N, box, particle_types=("A",), bond_types=(), **kw: _Snapshot(N, box, particle_types,
                                                                                          bond_types)


def _require_context():
    if context.current is None:
        raise RuntimeError("Cannot initialize without context.initialize() first")


def _read_snapshot(snapshot):
    _require_context()
    if context.current.system is not None:
        raise RuntimeError("Cannot initialize more than once")
    context.current.system = _SystemData(snapshot)
    return context.current.system


def _create_lattice(unitcell, n):
    """hoomd.init.create_lattice (hoomd/init.py:32-93): n unit cells per direction, box centred on the origin"""
    n = (n, n, n) if np.isscalar(n) else tuple(n)
    L = np.array([unitcell.a1[0] * n[0], unitcell.a2[1] * n[1], unitcell.a3[2] * n[2]])
    types_ = sorted(set(unitcell.type_name), key=unitcell.type_name.index)
    N = unitcell.N * n[0] * n[1] * n[2]
    snap = _Snapshot(N, _BoxDim(*L), types_)
    i, j, k = np.meshgrid(np.arange(n[0]), np.arange(n[1]), np.arange(n[2]), indexing="ij")
    cells = np.stack([i.ravel() * unitcell.a1[0], j.ravel() * unitcell.a2[1], k.ravel() * unitcell.a3[2]], axis=1)
    lo = -L / 2.0
    # SnapshotParticleData::replicate order (hoomd/ParticleData.cc:2892-2930): unit-cell particle outermost, then x, y, z
    half = np.array([unitcell.a1[0], unitcell.a2[1], unitcell.a3[2]]) / 2.0
    pos = (unitcell.position[:, None, :] + cells[None, :, :] + lo + half).reshape(-1, 3)
    snap.particles.position[:] = pos
    snap.particles.typeid[:] = np.repeat([types_.index(t) for t in unitcell.type_name], n[0] * n[1] * n[2])
    return _read_snapshot(snap)


def _gsd_snapshot(filename, frame=0):
    """hoomd.data.gsd_snapshot (hoomd/data.py, GSDReader): one frame of a GSD file as a snapshot"""
    from . import gsd_io
    fr = gsd_io.read_hoomd_frame(filename, frame)
    b = fr["box"]
    if b[3] != 0.0 or b[4] != 0.0 or b[5] != 0.0:
        raise RuntimeError("hoomd_b200: triclinic boxes are not supported (tilt factors in " + filename + ")")
    snap = _Snapshot(fr["N"], _BoxDim(float(b[0]), float(b[1]), float(b[2]), dimensions=fr["dimensions"]), fr["types"],
                     fr["bond_types"])
    q = snap.particles
    q.position[:], q.velocity[:], q.image[:] = fr["position"], fr["velocity"], fr["image"]
    q.typeid[:], q.mass[:], q.diameter[:] = fr["typeid"], fr["mass"], fr["diameter"]
    snap.bonds.resize(fr["bonds_N"])
    if fr["bonds_N"]:
        snap.bonds.group[:], snap.bonds.typeid[:] = fr["bonds_group"], fr["bonds_typeid"]
    snap._gsd_step = fr["step"]
    return snap


def _read_gsd(filename, restart=None, frame=0, time_step=None):
    """hoomd.init.read_gsd (hoomd/init.py:256-313): initialise from `filename`, or from `restart` (its frame 0) when that
    file exists; the time step comes from the file unless given"""
    import os
    _require_context()
    if context.current.system is not None:
        raise RuntimeError("Cannot initialize more than once")
    reading = filename
    if restart is not None and os.path.exists(restart):
        reading, frame = restart, 0
    snap = _gsd_snapshot(reading, frame)
    context.current.timestep = int(snap._gsd_step if time_step is None else time_step)
    return _read_snapshot(snap)


data.gsd_snapshot = _gsd_snapshot
init = types.ModuleType("hoomd.init")
init.read_snapshot = _read_snapshot
init.create_lattice = _create_lattice
init.read_gsd = _read_gsd

# ---- hoomd.group ---------------------------------------------------------------------------------------------------------
group = types.ModuleType("hoomd.group")


class _Group:
    """hoomd.group (hoomd/group.py): a named set of particle TAGS (ParticleGroup, hoomd/ParticleGroup.h); tags None = all"""

    def __init__(self, name, tags=None):
        _require_context()
        if context.current.system is None:
            raise RuntimeError("Cannot create a group before initialization")
        self.name = name
        self.tags = None if tags is None else np.unique(np.asarray(tags, np.uint32))
        if self.tags is not None and len(self.tags) and self.tags.max() >= context.current.system.N:
            raise RuntimeError("Error creating group: particle tag out of bounds")

    def __len__(self):
        return context.current.system.N if self.tags is None else len(self.tags)

    def is_all(self):
        return self.tags is None or len(self.tags) == context.current.system.N


def _group_type(type, name=None, update=False):
    """group.type(type): every particle of that type at creation (hoomd/group.py:329-372)"""
    sysd = context.current.system
    if type not in sysd.types:
        raise RuntimeError("Error creating group: type " + str(type) + " not found")
    sysd.take_snapshot()
    tid = sysd.pos4[:, 3].view(np.int32)
    return _Group(name or ("type " + type), np.nonzero(tid == sysd.types.index(type))[0])


group.all = lambda: _Group("all")
group.tag_list = lambda name, tags: _Group(name, tags)                                   # hoomd/group.py:297-327
group.tags = lambda tag_min, tag_max=None, name=None, update=False: _Group(              # hoomd/group.py:260-295
    name or "tags %d-%d" % (tag_min, tag_min if tag_max is None else tag_max),
    np.arange(tag_min, (tag_min if tag_max is None else tag_max) + 1))
group.type = _group_type

# ---- hoomd.update ---------------------------------------------------------------------------------------------------------

class _Scheduled:
    """period / phase bookkeeping of an analyzer or updater exactly as System::addAnalyzer / addUpdater and
    analyzer_item::shouldExecute keep it (hoomd/System.cc:69-95,212-243, hoomd/System.h:205-266): the first execution is
    start_step = ceil(cur / period) * period + phase (phase -1: the current step), every execution adds `period`."""

    def _schedule(self, period, phase):
        cur = context.current.timestep
        self.period, self.phase = int(period), int(phase)
        if self.phase >= 0:
            self.next_execute = (cur // self.period + (1 if cur % self.period else 0)) * self.period + self.phase
        else:
            self.next_execute = cur

    def should_execute(self, ts):
        """true on the scheduled step; maintains state like shouldExecute (call once per step)"""
        if ts == self.next_execute:
            self.next_execute += self.period
            return True
        return False

    def steps_until(self, ts):
        """steps from ts to the next execution (None when it lies in the past and will never fire)"""
        return self.next_execute - ts if self.next_execute > ts else None


update = types.ModuleType("hoomd.update")


class _Sort:
    """hoomd.update.sort: period 300 on the GPU (hoomd/update.py:258-263)"""

    def __init__(self):
        self.period, self.enabled = 300, True

    def set_params(self, grid=None):
        pass

    def set_period(self, period):
        self.period = int(period)
        if context.current is not None:
            context.current.engine_dirty = True

    def disable(self):
        self.enabled = False
        if context.current is not None:
            context.current.engine_dirty = True

    def enable(self):
        self.enabled = True


update.sort = _Sort

# ---- hoomd.variant (hoomd/variant.py, hoomd/Variant.cc) --------------------------------------------------------------------
variant = types.ModuleType("hoomd.variant")


class _VariantConst:
    def __init__(self, value):
        self.value = float(value)

    def getValue(self, timestep):
        return self.value


class _LinearInterp:
    """hoomd.variant.linear_interp: piecewise linear in the time step, constant outside the points (Variant.cc:38-91)"""

    def __init__(self, points, zero="now"):
        _require_context()
        if len(points) == 0:
            raise RuntimeError("Error creating variant: cannot create a linear_interp variant with 0 points")
        now = context.current.timestep
        if zero == "now":
            self.offset = now
        else:
            if zero < 0:
                raise RuntimeError("Error creating variant: cannot create a linear_interp variant with a negative zero")
            if zero > now:
                raise RuntimeError("Error creating variant: cannot create a linear_interp variant with a zero in the future")
            self.offset = int(zero)
        self.points = {}
        for t, v in points:
            if t < 0:
                raise RuntimeError("Error creating variant: negative timestep in a variant is not allowed")
            self.points[int(t)] = float(v)
        self.keys = sorted(self.points)

    def getValue(self, timestep):
        t = 0 if timestep < self.offset else timestep - self.offset
        k = self.keys
        if len(k) == 1 or t < k[0]:
            return self.points[k[0]]
        if t >= k[-1]:
            return self.points[k[-1]]
        import bisect
        b = bisect.bisect_right(k, t)
        ta, tb = k[b - 1], k[b]
        f = float(t - ta) / float(tb - ta)
        return (1.0 - f) * self.points[ta] + f * self.points[tb]


def _setup_variant_input(v):
    return v if hasattr(v, "getValue") else _VariantConst(v)


variant.linear_interp = _LinearInterp
variant._setup_variant_input = _setup_variant_input


class _BoxResize(_Scheduled):
    """hoomd.update.box_resize (hoomd/update.py:281-390, hoomd/BoxResizeUpdater.cc): every `period` steps the box is set to
    the variants' values and the particles are scaled with it (or wrapped into it).  Orthorhombic boxes only."""

    def __init__(self, Lx=None, Ly=None, Lz=None, xy=None, xz=None, yz=None, period=1, L=None, phase=0,
                 scale_particles=True):
        _require_context()
        ctx = context.current
        if ctx.system is None:
            raise RuntimeError("Cannot create box_resize before initialization")
        if L is not None:
            Lx = Ly = Lz = L
        if Lx is None and Ly is None and Lz is None and xy is None and xz is None and yz is None:
            raise RuntimeError("update.box_resize: one of Lx, Ly, Lz, xy, xz, yz must be set")
        for t in (xy, xz, yz):
            if t is not None and (hasattr(t, "getValue") or float(t) != 0.0):
                raise RuntimeError("hoomd_b200: triclinic boxes are out of scope (tilt factors must stay 0)")
        box = ctx.system.box
        self.Lx = _setup_variant_input(box.Lx if Lx is None else Lx)
        self.Ly = _setup_variant_input(box.Ly if Ly is None else Ly)
        self.Lz = _setup_variant_input(box.Lz if Lz is None else Lz)
        self.scale_particles = bool(scale_particles)
        self.enabled = True
        self.period = None if period is None else int(period)
        self.phase = int(phase)
        if period is None:
            self._apply(ctx.timestep)  # applied immediately, no periodic updates (hoomd/update.py:386-387)
        else:
            if self.period < 1:
                raise RuntimeError("update.box_resize: period must be a positive integer")
            self._schedule(self.period, self.phase)  # System::addUpdater
            ctx.box_resizers.append(self)

    def _apply(self, ts):
        ctx = context.current
        L3 = [np.float32(v.getValue(ts)) for v in (self.Lx, self.Ly, self.Lz)]  # Scalar Lx = m_Lx->getValue(timestep)
        if ctx.engine is None or ctx.engine_dirty:
            _configure_engine()
        ctx.engine.set_box(L3[0], L3[1], L3[2], self.scale_particles)
        ctx.system.box = _BoxDim(float(L3[0]), float(L3[1]), float(L3[2]))

    def disable(self):
        self.enabled = False  # the reference removes the updater from the system (hoomd/update.py:108-125)

    def enable(self):
        if not self.enabled and self.period is not None:
            self._schedule(self.period, self.phase)  # re-added with addUpdater (hoomd/update.py:127-145)
        self.enabled = True

    def set_period(self, period):
        self._schedule(int(period), self.phase)  # System::setUpdaterPeriod (hoomd/update.py:147-175)


update.box_resize = _BoxResize

# ---- hoomd.md -------------------------------------------------------------------------------------------------------------
md = types.ModuleType("hoomd.md")
md.nlist = types.ModuleType("hoomd.md.nlist")
md.pair = types.ModuleType("hoomd.md.pair")
md.integrate = types.ModuleType("hoomd.md.integrate")


class _Cell:
    """md.nlist.cell (hoomd/md/nlist.py:555)"""

    def __init__(self, r_buff=0.4, check_period=1, d_max=None, dist_check=True, name=None, deterministic=False):
        _require_context()
        if context.current.system is None:
            raise RuntimeError("Cannot create neighbor list before initialization")
        self.r_buff, self.check_period, self.dist_check = float(r_buff), int(check_period), bool(dist_check)
        self.exclusions = ["bond", "body"]  # the md.nlist default (hoomd/md/nlist.py:104-110)
        self.subscribers = []
        context.current.neighbor_lists.append(self)
        context.current.engine_dirty = True

    def set_params(self, r_buff=None, check_period=None, d_max=None, dist_check=None):
        if r_buff is not None:
            self.r_buff = float(r_buff)
        if check_period is not None:
            self.check_period = int(check_period)
        if dist_check is not None:
            self.dist_check = bool(dist_check)
        context.current.engine_dirty = True

    def reset_exclusions(self, exclusions=None):
        """nlist.reset_exclusions (hoomd/md/nlist.py:213-285): 'bond' and 'body' are honoured here; the angle / dihedral /
        1-3 / 1-4 kinds belong to force fields outside this path"""
        exclusions = list(exclusions or [])
        for e in exclusions:
            if e not in ("bond", "body", "constraint", "pair"):
                raise RuntimeError("hoomd_b200: exclusion type '%s' is not supported on this path" % e)
        self.exclusions = exclusions
        context.current.engine_dirty = True

    def query_update_period(self):
        eng = context.current.engine
        return None if eng is None else eng.stats()


md.nlist.cell = _Cell


class _Coeff:
    """pair.coeff (hoomd/md/pair.py:60-255): symmetric type-pair table with per-potential defaults"""

    def __init__(self):
        self.values = {}
        self.default_coeff = {}

    def set_default_coeff(self, name, value):
        self.default_coeff[name] = value

    def set(self, a, b, **coeffs):
        a = [a] if isinstance(a, str) else list(a)
        b = [b] if isinstance(b, str) else list(b)
        for ai in a:
            for bi in b:
                key = (ai, bi) if (ai, bi) in self.values or (bi, ai) not in self.values else (bi, ai)
                self.values.setdefault(key, {}).update(coeffs)
        if context.current is not None:
            context.current.engine_dirty = True

    def get(self, a, b, name):
        v = self.values.get((a, b), self.values.get((b, a)))
        if v is None:
            return None
        return v.get(name, self.default_coeff.get(name))

    def verify(self, types_, required):
        ok = True
        for i, a in enumerate(types_):
            for b in types_[i:]:
                if (a, b) not in self.values and (b, a) not in self.values:
                    sys.stderr.write("**ERROR**: Type pair (%s, %s) not found in pair coeff\n" % (a, b))
                    ok = False
                    continue
                for r in required:
                    if self.get(a, b, r) is None:
                        sys.stderr.write("**ERROR**: Type pair (%s, %s) is missing required coefficient %s\n" % (a, b, r))
                        ok = False
        return ok


class _Pair:
    """md.pair.pair (hoomd/md/pair.py:258-445): one PotentialPair<evaluator> on a neighbor list"""
    kind, required_coeffs, defaults = None, [], {}

    def __init__(self, r_cut, nlist, name=None):
        _require_context()
        if context.current.system is None:
            raise RuntimeError("Cannot create force before initialization")
        if any(isinstance(f, _Pair) and f.enabled for f in context.current.forces):
            raise RuntimeError("hoomd_b200: one enabled pair force per system in this round")
        self.global_r_cut, self.nlist, self.shift_mode = float(r_cut), nlist, 0
        self.pair_coeff = _Coeff()
        for k, v in self.defaults.items():
            self.pair_coeff.set_default_coeff(k, v)
        self.pair_coeff.set_default_coeff("r_cut", self.global_r_cut)
        self.pair_coeff.set_default_coeff("r_on", self.global_r_cut)
        self.required_coeffs = list(type(self).required_coeffs)
        self.enabled = True
        nlist.subscribers.append(self)
        context.current.forces.append(self)
        context.current.engine_dirty = True

    def set_params(self, mode=None):
        if mode is not None:
            if mode not in ("no_shift", "shift", "xplor"):
                raise RuntimeError("Error changing parameters in pair force")  # hoomd/md/pair.py:402-403
            self.shift_mode = {"no_shift": 0, "shift": 1, "xplor": 2}[mode]
            context.current.engine_dirty = True

    def tables(self, types_):
        """ntypes x ntypes (r_cut, epsilon, sigma, alpha, r_on); raises like update_coeffs (hoomd/md/pair.py:430-445)"""
        if not self.pair_coeff.verify(types_, self.required_coeffs):
            raise RuntimeError("Error updating pair coefficients")
        nt = len(types_)
        out = {k: np.zeros((nt, nt)) for k in ["r_cut", "r_on"] + self.required_coeffs}
        for i, a in enumerate(types_):
            for j, b in enumerate(types_):
                for k in out:
                    v = self.pair_coeff.get(a, b, k)
                    if v is False:  # r_cut = False disables the pair (hoomd/md/pair.py:330-345)
                        v = -1.0
                    out[k][i, j] = v
        return out

    def get_energy(self, group=None):
        """calcEnergySum of this ForceCompute alone (the pair component, not the net force)"""
        _ensure_forces(components=True)
        f, _ = context.current.engine.force_component(0)
        return float(f[:, 3].astype(np.float64).sum())


class _LJ(_Pair):
    """md.pair.lj (hoomd/md/pair.py:579-609)"""
    kind, required_coeffs, defaults = "lj", ["epsilon", "sigma", "alpha"], {"alpha": 1.0}


class _Gauss(_Pair):
    """md.pair.gauss (hoomd/md/pair.py:611-678)"""
    kind, required_coeffs = "gauss", ["epsilon", "sigma"]


class _Yukawa(_Pair):
    """md.pair.yukawa (hoomd/md/pair.py:806-869)"""
    kind, required_coeffs = "yukawa", ["epsilon", "kappa"]


class _Mie(_Pair):
    """md.pair.mie (hoomd/md/pair.py:2017-2093)"""
    kind, required_coeffs = "mie", ["epsilon", "sigma", "n", "m"]


class _ForceShiftedLJ(_Pair):
    """md.pair.force_shifted_lj (hoomd/md/pair.py:1692-1772)"""
    kind, required_coeffs, defaults = "force_shifted_lj", ["epsilon", "sigma", "alpha"], {"alpha": 1.0}


class _Morse(_Pair):
    """md.pair.morse (hoomd/md/pair.py:1213-1281)"""
    kind, required_coeffs = "morse", ["D0", "alpha", "r0"]


md.pair.lj = _LJ
md.pair.gauss = _Gauss
md.pair.yukawa = _Yukawa
md.pair.mie = _Mie
md.pair.force_shifted_lj = _ForceShiftedLJ
md.pair.morse = _Morse


class _DPDConservative(_Pair):
    """md.pair.dpd_conservative (hoomd/md/pair.py:1440-1517): the conservative part of the DPD force"""
    kind, required_coeffs = "dpd_conservative", ["A"]

    def set_params(self, coeff=None, mode=None):
        """dpd_conservative has no energy shift modes (hoomd/md/pair.py:1518-1522 raises unconditionally)"""
        raise RuntimeError("Not implemented for DPD Conservative")


md.pair.dpd_conservative = _DPDConservative


class _DPD(_Pair):
    """md.pair.dpd (hoomd/md/pair.py:1282-1438): conservative force + pairwise drag and random forces -- the DPD thermostat,
    to be used with md.integrate.nve.  kT must be a constant here (a variant would be evaluated per step by the caller)."""
    kind, required_coeffs = "dpd", ["A", "gamma"]

    def __init__(self, r_cut, nlist, kT, seed, name=None):
        super().__init__(r_cut, nlist, name)
        self.seed = int(seed)
        self.kT_variant = _setup_variant_input(kT)
        if not isinstance(self.kT_variant, _VariantConst):
            raise RuntimeError("hoomd_b200: pair.dpd takes a constant kT")

    def set_params(self, kT=None, mode=None):
        if mode is not None:
            raise RuntimeError("pair.dpd has no energy shift modes")
        if kT is not None:
            self.kT_variant = _setup_variant_input(kT)
            if not isinstance(self.kT_variant, _VariantConst):
                raise RuntimeError("hoomd_b200: pair.dpd takes a constant kT")
            context.current.engine_dirty = True

    @property
    def kT(self):
        return float(np.float32(self.kT_variant.getValue(context.current.timestep)))


md.pair.dpd = _DPD


class _BondCoeff:
    """bond.coeff (hoomd/md/bond.py:22-164): per-bond-type table"""

    def __init__(self):
        self.values = {}

    def set(self, type, **coeffs):
        for t in ([type] if isinstance(type, str) else list(type)):
            self.values.setdefault(t, {}).update(coeffs)
        if context.current is not None:
            context.current.engine_dirty = True

    def table(self, bond_types, required):
        out = {k: [] for k in required}
        for t in bond_types:
            if t not in self.values:
                sys.stderr.write("**ERROR**: Bond type %s not found in bond coeff\n" % t)
                raise RuntimeError("Error updating bond coefficients")
            for k in required:
                if k not in self.values[t]:
                    sys.stderr.write("**ERROR**: Bond type %s is missing required coefficient %s\n" % (t, k))
                    raise RuntimeError("Error updating bond coefficients")
                out[k].append(float(self.values[t][k]))
        return out


class _Bond:
    kind, required_coeffs = None, []

    def __init__(self, name=None):
        _require_context()
        if context.current.system is None:
            raise RuntimeError("Cannot create force before initialization")
        if len(context.current.system.bonds) == 0:
            sys.stderr.write("**ERROR**: No bonds are defined.\n")
            raise RuntimeError("Error creating bond forces")  # hoomd/md/bond.py:348-351
        if any(isinstance(f, _Bond) for f in context.current.forces):
            raise RuntimeError("hoomd_b200: one bond force per system in this round")
        self.bond_coeff = _BondCoeff()
        self.enabled = True
        context.current.forces.append(self)
        context.current.engine_dirty = True

    def get_energy(self, group=None):
        """calcEnergySum of the bond ForceCompute alone"""
        _ensure_forces(components=True)
        f, _ = context.current.engine.force_component(1)
        return float(f[:, 3].astype(np.float64).sum())


class _FENE(_Bond):
    """md.bond.fene (hoomd/md/bond.py:308-372)"""
    kind, required_coeffs = "fene", ["k", "r0", "epsilon", "sigma"]


class _Harmonic(_Bond):
    """md.bond.harmonic (hoomd/md/bond.py:236-296)"""
    kind, required_coeffs = "harmonic", ["k", "r0"]


md.bond = types.ModuleType("hoomd.md.bond")
md.bond.fene = _FENE
md.bond.harmonic = _Harmonic


class _ModeStandard:
    """md.integrate.mode_standard (hoomd/md/integrate.py:49)"""

    def __init__(self, dt, aniso=None):
        _require_context()
        self.dt = float(dt)
        context.current.integrator = self
        context.current.engine_dirty = True

    def set_params(self, dt=None, aniso=None):
        if dt is not None:
            self.dt = float(dt)
            context.current.engine_dirty = True


class _Method:
    def __init__(self, group_):
        _require_context()
        if group_ is None or not isinstance(group_, _Group):
            raise RuntimeError("integrate: a hoomd.group is required")
        self.group = group_
        context.current.integration_methods.append(self)
        context.current.engine_dirty = True
        self.enabled = True

    def disable(self):
        self.enabled = False
        context.current.engine_dirty = True

    def enable(self):
        self.enabled = True
        context.current.engine_dirty = True

    @property
    def kT(self):
        """set point at the current time step: kT may be a hoomd.variant (Scalar T = m_T->getValue(timestep))"""
        return float(np.float32(self.kT_variant.getValue(context.current.timestep)))

    @property
    def kT_is_variable(self):
        return not isinstance(getattr(self, "kT_variant", None), (_VariantConst, type(None)))


class _NVE(_Method):
    """md.integrate.nve (hoomd/md/integrate.py:729)"""

    def __init__(self, group, limit=None, zero_force=False):
        super().__init__(group)
        self.limit, self.zero_force = limit, bool(zero_force)

    def set_params(self, limit=None, zero_force=None):
        if limit is not None:
            self.limit = None if limit is False else limit
        if zero_force is not None:
            self.zero_force = bool(zero_force)
        context.current.engine_dirty = True


class _Langevin(_Method):
    """md.integrate.langevin (hoomd/md/integrate.py:844)"""

    def __init__(self, group, kT, seed, dscale=False, tally=False, noiseless_t=False, noiseless_r=False):
        super().__init__(group)
        # dscale = False, or lambda: gamma = lambda * diameter (use_lambda, hoomd/md/integrate.py:905-916)
        self.dscale = None if (dscale is False or dscale is None) else float(dscale)
        self.tally = bool(tally)
        self.kT_variant, self.seed, self.noiseless_t = _setup_variant_input(kT), int(seed), bool(noiseless_t)
        self.gamma = {}

    def set_params(self, kT=None, tally=None):
        if kT is not None:
            self.kT_variant = _setup_variant_input(kT)
            context.current.engine_dirty = True
        if tally is not None:
            self.tally = bool(tally)
            context.current.engine_dirty = True

    def set_gamma(self, a, gamma):
        types_ = context.current.system.types
        if isinstance(a, str):
            if a not in types_:
                raise RuntimeError("Trying to set gamma for a non existent type! " + str(a))
            self.gamma[a] = float(gamma)
        else:
            self.gamma[types_[a]] = float(gamma)
        context.current.engine_dirty = True


class _NVT(_Method):
    """md.integrate.nvt (hoomd/md/integrate.py:163-260): Nose-Hoover thermostat, MTK equations"""

    def __init__(self, group, kT, tau):
        super().__init__(group)
        self.kT_variant, self.tau = _setup_variant_input(kT), float(tau)

    def set_params(self, kT=None, tau=None):
        if kT is not None:
            self.kT_variant = _setup_variant_input(kT)
        if tau is not None:
            self.tau = float(tau)
        context.current.engine_dirty = True


class _NPT(_Method):
    """md.integrate.npt (hoomd/md/integrate.py:262-620): Martyna-Tobias-Klein barostat + Nose-Hoover thermostat
    (TwoStepNPTMTK).  Orthorhombic boxes: the tilt degrees of freedom xy / xz / yz (and all=True) raise."""

    def __init__(self, group, kT=None, tau=None, S=None, P=None, tauP=None, couple="xyz", x=True, y=True, z=True,
                 xy=False, xz=False, yz=False, all=False, nph=False, rescale_all=None, gamma=None):
        if kT is None or tau is None:
            if nph is False:
                raise RuntimeError("Error setting up NPT integration.")  # integrate.npt: Need temperature T and tau
            kT, tau = 1.0, 1.0  # "use dummy values" (hoomd/md/integrate.py:414-417)
        if tauP is None:
            raise RuntimeError("Error setting up NPT integration.")      # integrate.npt: Need barostat time scale tauP
        if P is not None:
            S = [P, P, P, 0, 0, 0]
        elif S is None or len(S) != 6:
            raise RuntimeError("Unrecognized stress tensor form")
        if couple not in ("none", "xy", "xz", "yz", "xyz"):
            raise RuntimeError("Error setting up NPT integration.")      # Invalid coupling mode
        if xy or xz or yz or all:
            raise RuntimeError("hoomd_b200: integrate.npt with tilt degrees of freedom needs triclinic boxes")
        if any(float(v) != float(S[0]) for v in S[1:3]) or any(float(v) != 0.0 for v in S[3:]):
            raise RuntimeError("hoomd_b200: integrate.npt supports an isotropic pressure set point (S = [P, P, P, 0, 0, 0])")
        super().__init__(group)
        self.kT_variant, self.tau, self.tauP = _setup_variant_input(kT), float(tau), float(tauP)
        self.P_variant = _setup_variant_input(S[0])
        if not isinstance(self.P_variant, _VariantConst):
            raise RuntimeError("hoomd_b200: integrate.npt takes a constant pressure set point (no pressure ramps yet)")
        self.couple, self.x, self.y, self.z = couple, bool(x), bool(y), bool(z)
        self.nph, self.rescale_all = bool(nph), bool(rescale_all) if rescale_all is not None else False
        self.gamma = 0.0 if gamma is None else float(gamma)
        self._state = None

    def set_params(self, kT=None, tau=None, S=None, P=None, tauP=None, rescale_all=None, gamma=None):
        if kT is not None:
            self.kT_variant = _setup_variant_input(kT)
        if tau is not None:
            self.tau = float(tau)
        if P is not None:
            self.P_variant = _setup_variant_input(P)
        elif S is not None:
            self.P_variant = _setup_variant_input(S[0])
        if tauP is not None:
            self.tauP = float(tauP)
        if rescale_all is not None:
            self.rescale_all = bool(rescale_all)
        if gamma is not None:
            self.gamma = float(gamma)
        context.current.engine_dirty = True

    @property
    def P(self):
        return float(np.float32(self.P_variant.getValue(context.current.timestep)))


md.integrate.mode_standard = _ModeStandard
md.integrate.nvt = _NVT
md.integrate.npt = _NPT
md.integrate.nve = _NVE
md.integrate.langevin = _Langevin

# ---- hoomd.analyze.log (in-memory subset: the thermodynamic quantities of ComputeThermo) ---------------------------------
analyze = types.ModuleType("hoomd.analyze")


class _Log(_Scheduled):
    QUANTITIES = {"temperature": "T", "kinetic_energy": "ke", "potential_energy": "pe", "pressure": "P"}

    def __init__(self, filename=None, quantities=(), period=1, header_prefix="", overwrite=False, phase=0, **kw):
        _require_context()
        import re
        for q in quantities:
            # per-force energies are logged under "<force name>_energy": pair_<kind>_energy (hoomd/md/PotentialPair.h:200-215),
            # bond_<kind>_energy (hoomd/md/PotentialBond.h:118-133); whether that force exists is checked when it is queried
            if q not in self.QUANTITIES and q not in ("nvt_mtk_reservoir_energy_all", "langevin_reservoir_energy_all",
                                                      "npt_thermostat_energy", "npt_barostat_energy", "volume", "lx", "ly",
                                                      "lz") \
                    and not re.fullmatch(r"(pair|bond)_[a-z_]+_energy", q):
                raise RuntimeError("Unknown log quantity " + q)
        self.quantities, self.filename, self.rows = list(quantities), filename, []
        self.header_prefix, self.overwrite, self._opened = header_prefix, bool(overwrite), False
        self.enabled = True
        self._schedule(int(period), int(phase))  # System::addAnalyzer
        context.current.loggers.append(self)
        # the virial is only evaluated when a logger asks for the pressure (pdata flags, hoomd/System.cc:855): a new
        # pressure logger re-configures the engine, any other logger only reads what the engine computes anyway
        if "pressure" in self.quantities:
            context.current.engine_dirty = True

    def set_period(self, period):
        self._schedule(int(period), self.phase)  # System::setAnalyzerPeriod

    def disable(self):
        self.enabled = False

    def enable(self):
        if not self.enabled:
            self._schedule(self.period, self.phase)
        self.enabled = True

    def _write(self, row):
        """plain-text log file in the reference's format (hoomd/LogPlainTXT.cc:100-195): tab-delimited, a header line for a
        new file, values with 10 significant digits; an existing file is appended to unless overwrite is set"""
        if self.filename is None:
            return
        import os
        if not self._opened:
            new = self.overwrite or not os.path.exists(self.filename)
            with open(self.filename, "w" if new else "a") as f:
                if new:
                    f.write(self.header_prefix + "\t".join(["timestep"] + self.quantities) + "\n")
            self._opened = True
        with open(self.filename, "a") as f:
            f.write("\t".join(["%d" % row[0]] + ["%.10g" % v for v in row[1:]]) + "\n")

    def query(self, quantity):
        if quantity == "langevin_reservoir_energy_all":  # md.integrate.langevin(tally=True) (hoomd/md/integrate.py:918-921)
            return context.current.engine.langevin_reservoir_energy()
        if quantity == "nvt_mtk_reservoir_energy_all":  # provided by md.integrate.nvt (log name + "_" + group name)
            return context.current.engine.nvt_reservoir_energy()
        if quantity in ("npt_thermostat_energy", "npt_barostat_energy"):  # TwoStepNPTMTK::getLogValue
            return float(context.current.engine.npt_state()[10 if quantity == "npt_thermostat_energy" else 11])
        if quantity in ("volume", "lx", "ly", "lz"):  # System-provided box quantities (hoomd/System.cc / analyze.log docs)
            L3 = [float(v) for v in context.current.engine.box_lengths()]
            return L3[0] * L3[1] * L3[2] if quantity == "volume" else L3["xyz".index(quantity[1])]
        ctx = context.current
        if quantity not in self.QUANTITIES:
            # the energy of ONE ForceCompute (ForceCompute::calcEnergySum, hoomd/ForceCompute.cc:82-102), not the total
            kind, name = quantity.split("_", 1)[0], quantity[:-len("_energy")].split("_", 1)[1]
            cls = _Pair if kind == "pair" else _Bond
            active = [f for f in ctx.forces if isinstance(f, cls) and f.enabled and f.kind == name]
            if not active:
                raise RuntimeError("Unknown log quantity " + quantity + ": no such force is defined")
            return active[0].get_energy()
        return float(ctx.engine.thermo()[self.QUANTITIES[quantity]])


analyze.log = _Log

# ---- hoomd.dump.gsd ----------------------------------------------------------------------------------------------------------
dump = types.ModuleType("hoomd.dump")


class _GSD(_Scheduled):
    """hoomd.dump.gsd (hoomd/dump.py:501-660, GSDDumpWriter): trajectory / restart frames of the HOOMD GSD schema written from
    the device-resident state (one download per frame).  period=None writes one frame now; truncate=True keeps rewriting
    frame 0 (restart files); `dynamic` / `static` select the chunk categories written after frame 0."""

    def __init__(self, filename, period, group, overwrite=False, truncate=False, phase=0, time_step=None, static=None,
                 dynamic=None):
        from . import gsd_io
        _require_context()
        if static is not None and dynamic is not None:
            raise ValueError("Cannot specify both static and dynamic arguments")
        if group is None or not isinstance(group, _Group):
            raise RuntimeError("dump.gsd: a hoomd.group is required")
        dyn = ["property"]
        if static is not None:
            dyn = [c for c in gsd_io.CATEGORIES if c not in static]
        if dynamic is not None:
            dyn = ["property"] + list(dynamic)
        self.filename, self.group, self.period, self.phase = filename, group, period, int(phase)
        self._writer_args = dict(overwrite=bool(overwrite), truncate=bool(truncate), dynamic=tuple(dyn))
        self._writer = None
        self.enabled = True
        ctx = context.current
        if period is not None:
            if int(period) < 1:
                raise RuntimeError("dump.gsd: period must be a positive integer")
            self._schedule(int(period), int(phase))  # System::addAnalyzer
            ctx.dumpers.append(self)
        else:
            self._analyze(ctx.timestep if time_step is None else int(time_step))

    def _analyze(self, ts):
        from . import gsd_io
        ctx = context.current
        if self._writer is None:  # GSDDumpWriter::initFileIO on the first frame
            self._writer = gsd_io.HoomdGSDWriter(self.filename, application="HOOMD-blue 2.3.0 (hoomd_b200)",
                                                 **self._writer_args)
        snap = ctx.system.take_snapshot()
        q = snap.particles
        tags = np.arange(q.N) if self.group.is_all() else np.asarray(self.group.tags, np.int64)
        frame = dict(position=q.position[tags], velocity=q.velocity[tags], image=q.image[tags], typeid=q.typeid[tags],
                     mass=q.mass[tags], diameter=q.diameter[tags], types=list(q.types))
        if self.group.is_all():  # topology is only meaningful for the all group (GSDDumpWriter.cc:245)
            frame.update(bonds_group=snap.bonds.group, bonds_typeid=snap.bonds.typeid, bond_types=list(snap.bonds.types))
        box = ctx.system.box
        self._writer.write_frame(ts, (box.Lx, box.Ly, box.Lz, 0.0, 0.0, 0.0), frame, dimensions=box.dimensions)

    def write_restart(self):
        """one frame at the current time step (restart files with truncate=True at the end of a job)"""
        self._analyze(context.current.timestep if context.current.engine is None else context.current.engine.timestep)

    def set_period(self, period):
        self._schedule(int(period), self.phase)

    def disable(self):
        self.enabled = False

    def enable(self):
        if not self.enabled and self.period is not None:
            self._schedule(self.period, self.phase)
        self.enabled = True


dump.gsd = _GSD


# ---- hoomd.run ------------------------------------------------------------------------------------------------------------
def _configure_engine():
    """(Re)build the device-resident system from the recorded commands: what System::run's prepRun / update_rcut /
    update_coeffs do in the reference before the first step."""
    from .engine import System
    ctx = context.current
    sysd = ctx.system
    if ctx.integrator is None:
        raise RuntimeError("Cannot run without an integrator")  # the reference warns and integrates nothing
    methods = [m for m in ctx.integration_methods if m.enabled]
    if len(methods) != 1:
        raise RuntimeError("hoomd_b200: exactly one enabled integration method is supported, got %d" % len(methods))
    ljs = [f for f in ctx.forces if isinstance(f, _Pair) and f.enabled]
    if len(ljs) != 1:
        raise RuntimeError("hoomd_b200: exactly one md.pair force is supported, got %d" % len(ljs))
    lj = ljs[0]
    accel = None
    if ctx.engine is not None:  # carry the state over
        # accelerations too: for Langevin the stored a(t) contains the previous step's random and drag force, which step one
        # of the next run must see (prepRun only computes a = F/m when no acceleration is set, IntegratorTwoStep.cc:379-392)
        accel = sysd.take_snapshot().particles.acceleration.astype(np.float32)
        for m in methods:  # IntegratorVariables live in the SystemDefinition, not in the method (hoomd/IntegratorData.h)
            if isinstance(m, _NVT) and getattr(ctx.engine, "_has_nvt", False):
                m._state = tuple(ctx.engine.nvt_state())
            if isinstance(m, _NPT) and getattr(ctx.engine, "_has_npt", False):
                m._state = ctx.engine.npt_state()[:10].copy()
        ctx.engine.close()
    eng = System(sysd.pos4, (sysd.box.Lx, sysd.box.Ly, sysd.box.Lz), ntypes=len(sysd.types), vel4=sysd.vel4,
                 image=sysd.image)
    t = lj.tables(sysd.types)
    need_virial = any("pressure" in lg.quantities for lg in ctx.loggers) or any(isinstance(m, _NPT) for m in methods)
    if lj.kind == "lj":
        eng.setup_lj(r_cut=t["r_cut"], epsilon=t["epsilon"], sigma=t["sigma"], alpha=t["alpha"], r_buff=lj.nlist.r_buff,
                     check_period=lj.nlist.check_period, shift_mode=lj.shift_mode,
                     r_on=t["r_on"] if lj.shift_mode == 2 else None)
    else:
        eng.setup_pair(lj.kind, r_cut=t["r_cut"], r_buff=lj.nlist.r_buff, check_period=lj.nlist.check_period,
                       shift_mode=lj.shift_mode, r_on=t["r_on"] if lj.shift_mode == 2 else None,
                       **{k: t[k] for k in lj.required_coeffs})
    if isinstance(lj, _DPD):
        eng.set_dpd_thermostat(lj.kT, lj.seed)
    bonds = [f for f in ctx.forces if isinstance(f, _Bond) and f.enabled]
    if len(sysd.bond_group):
        eng.set_bonds(sysd.bond_group, sysd.bond_typeid, n_bond_types=len(sysd.bond_types))
        if bonds:
            t = bonds[0].bond_coeff.table(sysd.bond_types, bonds[0].required_coeffs)
            if bonds[0].kind == "fene":
                eng.setup_fene(k=t["k"], r0=t["r0"], epsilon=t["epsilon"], sigma=t["sigma"])
            else:
                eng.setup_harmonic(k=t["k"], r0=t["r0"])
        eng.exclusions_from_bonds("bond" in lj.nlist.exclusions)
    eng.set_dist_check(lj.nlist.dist_check)
    eng.set_virial(need_virial)
    eng.setup_integrator(ctx.integrator.dt)
    m = methods[0]
    if isinstance(m, _NVE):
        eng.add_nve(limit=m.limit, zero_force=m.zero_force)
    elif isinstance(m, _NVT):
        eng.add_nvt(m.kT, m.tau)
        eng._has_nvt = True
        if getattr(m, "_state", None) is not None:
            eng.set_nvt_state(*m._state)
    elif isinstance(m, _NPT):
        eng.add_npt(m.kT, m.tau, m.P, m.tauP, couple=m.couple, x=m.x, y=m.y, z=m.z, nph=m.nph,
                    rescale_all=m.rescale_all, gamma=m.gamma)
        eng._has_npt = True
        if m._state is not None:
            eng.set_npt_state(m._state)
    else:
        gamma = [m.gamma.get(tn, 1.0) for tn in sysd.types]
        eng.add_langevin(m.kT, m.seed, gamma=gamma, noiseless_t=m.noiseless_t, dscale=m.dscale, tally=m.tally)
        if m.dscale is not None:
            eng.set_diameters(sysd.diameter)
    if not m.group.is_all():
        eng.set_group(m.group.tags)
    # the sorter is an updater created at step 0 with phase 0: it runs on multiples of its period (System::addUpdater), also
    # after a re-configuration in the middle of a job.  b200_sys_set_sorter schedules from the engine's current time step.
    sp = ctx.sorter.period if ctx.sorter.enabled else 0
    eng.set_timestep((ctx.timestep + sp - 1) // sp * sp if sp else ctx.timestep)
    eng.set_sorter(sp)
    eng.set_timestep(ctx.timestep)
    if accel is not None:
        eng.set_accel(accel)
    ctx.engine, ctx.engine_dirty, ctx.forces_stale = eng, False, True


def _ensure_forces(components=False):
    """The reference computes the forces in prepRun before the first step of a run, so analyzers at that step (and
    force.get_energy right after a run) see real values; here the first force evaluation of a freshly configured engine is
    made explicit.  components: the per-ForceCompute arrays are wanted -- the single-kernel MD step only keeps the NET
    force, so they are re-evaluated at the current step (same neighbor list: NeighborList::compute is memoised per step)."""
    ctx = context.current
    if ctx.engine is None or ctx.engine_dirty:
        _configure_engine()
    if getattr(ctx, "forces_stale", False) or components:
        ctx.engine.compute_forces(ctx.engine.timestep)
        ctx.forces_stale = False


def run(tsteps, profile=False, limit_hours=None, limit_multiple=1, callback_period=0, callback=None, quiet=False):
    """hoomd.run: advance tsteps steps; logged quantities are sampled every `period` steps (hoomd/__init__.py run)"""
    _require_context()
    ctx = context.current
    if ctx.system is None:
        raise RuntimeError("Cannot run before initialization")
    if ctx.engine is None or ctx.engine_dirty:
        _configure_engine()
    eng = ctx.engine
    tsteps = int(tsteps)
    cb_period = int(callback_period) if callback else 0
    done = 0
    while done < tsteps:
        ts = eng.timestep
        # System::run order within a step (hoomd/System.cc:552-610): callback, analyzers, updaters, integrator
        if callback and cb_period > 0 and ts % cb_period == 0:
            rv = callback(ts)
            if rv is not None and int(rv) < 0:
                break  # a negative return value ends the run
        for lg in ctx.loggers:
            if lg.enabled and lg.should_execute(ts):
                _ensure_forces()
                eng = ctx.engine
                lg.rows.append([ts] + [lg.query(q) for q in lg.quantities])
                lg._write(lg.rows[-1])
        for dp in ctx.dumpers:
            if dp.enabled and dp.should_execute(ts):
                dp._analyze(ts)
        for br in ctx.box_resizers:
            if br.enabled and br.should_execute(ts):
                br._apply(ts)
                eng = ctx.engine
        ramp = [m for m in ctx.integration_methods if m.enabled and m.kT_is_variable]
        for m in ramp:
            eng.set_kT(np.float32(m.kT_variant.getValue(ts)))
        # advance to the next scheduled event
        nxt = tsteps - done
        if ramp:  # a temperature variant is evaluated at every step, like m_T->getValue(timestep) in the reference
            nxt = 1
        if cb_period > 0:
            nxt = min(nxt, cb_period - ts % cb_period)
        for item in list(ctx.loggers) + list(ctx.dumpers) + list(ctx.box_resizers):
            if item.enabled:
                n = item.steps_until(ts)
                if n is not None:
                    nxt = min(nxt, n)
        eng.run(nxt)
        ctx.forces_stale = False
        done += nxt
        if getattr(eng, "_has_npt", False):  # the barostat moves the box every step (ParticleData::setGlobalBox)
            L3 = [float(v) for v in eng.box_lengths()]
            ctx.system.box = _BoxDim(L3[0], L3[1], L3[2])
    ctx.timestep = eng.timestep
    if callback and cb_period == 0:
        callback(ctx.timestep)  # callback_period 0: called once at the end of the run (hoomd/System.cc:617-621)


def get_step():
    return context.current.timestep if context.current else 0


context = _ContextModule()


def install():
    """register this module as `hoomd` (+ submodules) so that unmodified job scripts import it"""
    me = sys.modules[__name__]
    sys.modules["hoomd"] = me
    for name, mod in (("context", context), ("init", init), ("lattice", lattice), ("data", data), ("group", group),
                      ("update", update), ("variant", variant), ("md", md), ("analyze", analyze), ("dump", dump)):
        sys.modules["hoomd." + name] = mod
    for name, mod in (("nlist", md.nlist), ("pair", md.pair), ("integrate", md.integrate), ("bond", md.bond)):
        sys.modules["hoomd.md." + name] = mod
    return me
