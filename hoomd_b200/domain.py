"""Spatial domain decomposition over the GPUs of one box: the host-side half.

`find_decomposition` / `local_bounds` / `place_particles` restate hoomd/DomainDecomposition.cc:323-372, 531-557, 566-613
(pure numpy: they run on the CPU and are what the world_size-2 gloo tests cover).  `DomainSystem` builds one
`hoomd_b200.System` per rank on its sub-domain and attaches the peers: every rank exports ONE CUDA IPC handle, the
handles are all-gathered with torch.distributed (the only collective the data path ever issues from the host), and from
then on ghosts, migrants and barrier flags move as plain stores over NVLink from inside the step kernels.
"""
import ctypes as C

import numpy as np

from . import _lib
from .engine import System


def find_decomposition(L, nranks):
    """(nx, ny, nz) with nx*ny*nz == nranks minimising the inter-domain surface (DomainDecomposition.cc:323-372)."""
    Lx, Ly, Lz = (float(v) for v in L)
    best, best_area = (1, 1, nranks), Lx * Ly * (nranks - 1)
    found = False
    for nx in range(1, nranks + 1):
        for ny in range(1, nranks // nx + 1):
            for nz in range(1, nranks // (nx * ny) + 1):
                if nx * ny * nz != nranks:
                    continue
                area = Lx * Ly * (nz - 1) + Lx * Lz * (ny - 1) + Ly * Lz * (nx - 1)
                if area < best_area or not found:
                    best, best_area, found = (nx, ny, nz), area, True
    return best


def rank_to_grid(rank, grid):
    """rank = (iz*ny + iy)*nx + ix"""
    nx, ny, nz = grid
    return rank % nx, (rank // nx) % ny, rank // (nx * ny)


def local_bounds(L, grid, rank):
    """(lo[3], hi[3], periodic[3]) of the rank's box in FP32: uniform cut planes, non-periodic along split axes
    (DomainDecomposition::calculateLocalBox, DomainDecomposition.cc:531-557).  Same arithmetic as csrc/engine_mg.inl."""
    pos = rank_to_grid(rank, grid)
    lo, hi, per = np.zeros(3, np.float32), np.zeros(3, np.float32), np.zeros(3, np.uint8)
    for a in range(3):
        La = np.float32(L[a])
        glo = -(La / np.float32(2.0))
        ghi = La / np.float32(2.0)
        if grid[a] > 1:
            lo[a] = glo + La * (np.float32(pos[a]) / np.float32(grid[a]))
            hi[a] = ghi if pos[a] + 1 == grid[a] else glo + La * (np.float32(pos[a] + 1) / np.float32(grid[a]))
            per[a] = 0
        else:
            lo[a], hi[a], per[a] = glo, ghi, 1
    return lo, hi, per


def place_particles(pos, L, grid):
    """owner rank of every particle (DomainDecomposition::placeParticle): lo <= x < hi of exactly one domain"""
    pos = np.asarray(pos, np.float32)
    owner = np.zeros(len(pos), np.int64)
    idx = []
    for a in range(3):
        La = np.float32(L[a])
        glo = -(La / np.float32(2.0))
        if grid[a] > 1:
            cuts = np.array([glo + La * (np.float32(k) / np.float32(grid[a])) for k in range(1, grid[a])], np.float32)
            idx.append(np.searchsorted(cuts, pos[:, a], side="right"))
        else:
            idx.append(np.zeros(len(pos), np.int64))
    owner = (idx[2] * grid[1] + idx[1]) * grid[0] + idx[0]
    return owner


class DomainSystem(System):
    """One rank of a domain-decomposed system.  Every rank passes the SAME global snapshot (synthetic, generated
    deterministically on each rank) and keeps the particles it owns; tags are the global particle indices."""

    def __init__(self, pos4, box, r_ghost, vel4=None, image=None, grid=None, rank=None, world=None, device=None,
                 ncap=0, capacity_factor=0.0, connect=True, ntypes=1):
        import torch.distributed as dist
        L = _lib.lib()
        self.world = dist.get_world_size() if world is None else int(world)
        self.rank = dist.get_rank() if rank is None else int(rank)
        pos4 = np.ascontiguousarray(pos4, np.float32)
        self.N_global = int(pos4.shape[0])
        self.ntypes = int(ntypes)
        self.n_bond_types = 0
        self.grid = tuple(grid) if grid is not None else find_decomposition(box, self.world)
        self.global_L = tuple(float(v) for v in box)
        owner = place_particles(pos4[:, :3], box, self.grid)
        mine = np.nonzero(owner == self.rank)[0].astype(np.uint32)
        self.N = int(len(mine))
        p = _lib.MgParams()
        p.world, p.rank = self.world, self.rank
        p.grid = (C.c_uint * 3)(*self.grid)
        p.global_box = _lib.make_box(*self.global_L)
        p.N_global, p.N_local, p.Ncap = self.N_global, self.N, int(ncap)
        p.r_ghost, p.capacity_factor, p.ntypes = float(r_ghost), float(capacity_factor), int(ntypes)
        p.device = int(self.rank if device is None else device)
        self.h = C.c_void_p(L.b200_sys_create_mg(C.byref(p)))
        self._check(0)
        self.box = _lib.Box()
        handle = (C.c_ubyte * 64)()
        caps = (C.c_uint * 3)()
        self._check(L.b200_sys_mg_export(self.h, handle, C.byref(self.box), caps))
        self.Ncap, self.Gcap, self.mig_cap = int(caps[0]), int(caps[1]), int(caps[2])
        self._handle = bytes(handle)
        self.set_particles(pos4[mine], None if vel4 is None else np.asarray(vel4, np.float32)[mine],
                           None if image is None else np.asarray(image, np.int32)[mine])
        self._check(L.b200_sys_set_tags(self.h, mine.ctypes.data_as(C.c_void_p)))
        if connect:
            self.connect()

    def connect(self):
        """all-gather the IPC handles and attach the peers"""
        import torch.distributed as dist
        handles = [None] * self.world
        dist.all_gather_object(handles, self._handle)
        blob = b"".join(handles)
        buf = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
        self._check(_lib.lib().b200_sys_mg_connect(self.h, buf))
        dist.barrier()

    def _sync_n(self):
        self.N = int(_lib.lib().b200_sys_get_n(self.h))
        return self.N

    def particles(self):
        self._sync_n()
        return super().particles()

    def forces(self, virial=False):
        self._sync_n()
        return super().forces(virial)

    def run(self, nsteps):
        out = super().run(nsteps)
        self._sync_n()
        return out

    def mg_stats(self):
        out = (C.c_ulonglong * 8)()
        self._check(_lib.lib().b200_sys_mg_stats(self.h, out))
        return dict(N=int(out[0]), Ncap=int(out[1]), Gcap=int(out[2]), max_N=int(out[3]), max_migrants=int(out[4]),
                    max_ghosts_dir=int(out[5]), ghosts_sent=int(out[6]), step_barriers=int(out[7]))

    def mg_timing(self):
        """device-measured barrier waits: per kind (step, migrate, ghost plan, overflow vote) total microseconds, count and
        mean microseconds per barrier; boundary / total chunks of the step kernel"""
        out = (C.c_ulonglong * 10)()
        self._check(_lib.lib().b200_sys_mg_timing(self.h, out))
        names = ("step", "migrate", "ghost_plan", "overflow_vote")
        res = {}
        for k, n in enumerate(names):
            ns, cnt = int(out[2 * k]), int(out[2 * k + 1])
            res[n] = dict(total_us=ns * 1e-3, count=cnt, mean_us=(ns * 1e-3 / cnt) if cnt else 0.0)
        res["boundary_chunks"], res["chunks"] = int(out[8]), int(out[9])
        return res

    def ghosts(self):
        pos = np.empty((self.Gcap, 4), np.float32)
        tag = np.empty(self.Gcap, np.uint32)
        self._check(_lib.lib().b200_sys_mg_get_ghosts(self.h, pos.ctypes.data_as(C.c_void_p),
                                                      tag.ctypes.data_as(C.c_void_p)))
        ok = pos[:, 0] < 1e29
        return pos[ok], tag[ok]

    def gather(self, keys=("pos", "vel", "image")):
        """global arrays ordered by tag on every rank (validation / snapshots; not a hot path)"""
        import torch.distributed as dist
        p = self.particles()
        parts = [None] * self.world
        dist.all_gather_object(parts, {k: p[k] for k in tuple(keys) + ("tag",)})
        tag = np.concatenate([q["tag"] for q in parts])
        order = np.argsort(tag, kind="stable")
        out = {k: np.concatenate([q[k] for q in parts])[order] for k in keys}
        out["tag"] = tag[order]
        return out
