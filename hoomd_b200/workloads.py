"""Synthetic initial states for the BASELINE.json configurations.  Pure numpy, deterministic."""
import numpy as np


def sc_lattice(n_side, a, kT=None, seed=42, ntypes=1, type_fractions=None):
    """n_side^3 simple-cubic lattice with spacing a centred in a cubic box L = n_side*a.

    Positions follow hoomd.init.create_lattice(unitcell=sc(a), n) (hoomd/init.py:32-93): cell corners shifted so the
    box is [-L/2, L/2).  Velocities: Gaussian kT (mass 1) from numpy default_rng(seed), net momentum removed.
    Returns (pos4 float32 [N,4] with the type bits in .w, vel4 float32 [N,4] with mass in .w, L).
    """
    N = n_side ** 3
    L = np.float32(n_side * a)
    g = (np.arange(n_side, dtype=np.float64) + 0.5) * a - n_side * a / 2.0
    x, y, z = np.meshgrid(g, g, g, indexing="ij")
    pos = np.zeros((N, 4), np.float32)
    pos[:, 0] = x.ravel()
    pos[:, 1] = y.ravel()
    pos[:, 2] = z.ravel()
    types = np.zeros(N, np.int32)
    rng = np.random.default_rng(seed)
    if ntypes > 1:
        fr = np.asarray(type_fractions, np.float64)
        counts = np.floor(fr * N).astype(int)
        counts[0] += N - counts.sum()
        types = np.repeat(np.arange(ntypes, dtype=np.int32), counts)
        types = types[rng.permutation(N)]
    pos[:, 3] = types.view(np.float32)
    vel = np.zeros((N, 4), np.float32)
    vel[:, 3] = 1.0
    if kT is not None:
        v = rng.normal(0.0, np.sqrt(kT), size=(N, 3))
        v -= v.mean(axis=0)
        vel[:, :3] = v
    return pos, vel, float(L)


def fcc_lattice(n_cells, rho, kT=None, seed=42, ntypes=1, type_fractions=None):
    """4 * n_cells^3 particles on a face-centred cubic lattice at number density rho in a cubic box (hoomd.lattice.fcc +
    hoomd.init.create_lattice, hoomd/lattice.py / hoomd/init.py:32-93).  n_cells = 40 gives BASELINE's 256,000 particles.
    Types, velocities and the return value as sc_lattice."""
    N = 4 * n_cells ** 3
    a = (4.0 / rho) ** (1.0 / 3.0)
    L = np.float32(n_cells * a)
    g = np.arange(n_cells, dtype=np.float64) * a - n_cells * a / 2.0
    x, y, z = np.meshgrid(g, g, g, indexing="ij")
    corner = np.stack([x.ravel(), y.ravel(), z.ravel()], axis=1)
    basis = np.array([[0.25, 0.25, 0.25], [0.25, 0.75, 0.75], [0.75, 0.25, 0.75], [0.75, 0.75, 0.25]]) * a
    pos = np.zeros((N, 4), np.float32)
    pos[:, :3] = (corner[:, None, :] + basis[None, :, :]).reshape(N, 3)
    types = np.zeros(N, np.int32)
    rng = np.random.default_rng(seed)
    if ntypes > 1:
        fr = np.asarray(type_fractions, np.float64)
        counts = np.floor(fr * N).astype(int)
        counts[0] += N - counts.sum()
        types = np.repeat(np.arange(ntypes, dtype=np.int32), counts)
        types = types[rng.permutation(N)]
    pos[:, 3] = types.view(np.float32)
    vel = np.zeros((N, 4), np.float32)
    vel[:, 3] = 1.0
    if kT is not None:
        v = rng.normal(0.0, np.sqrt(kT), size=(N, 3))
        v -= v.mean(axis=0)
        vel[:, :3] = v
    return pos, vel, float(L)


def lj_liquid(N, rho=0.2, kT=1.2, seed=42):
    """LJ liquid initial lattice: N must be a cube; L = (N/rho)^(1/3)."""
    n_side = int(round(N ** (1.0 / 3.0)))
    assert n_side ** 3 == N, "N must be a perfect cube"
    a = (1.0 / rho) ** (1.0 / 3.0)
    return sc_lattice(n_side, a, kT=kT, seed=seed)


def random_system(N, L, seed=0, ntypes=1, min_dist=0.0):
    """Uniform random positions in [-L/2, L/2)^3 (the protocol of hoomd/md/test/test_neighborlist.cc:873-937)."""
    rng = np.random.default_rng(seed)
    pos = np.zeros((N, 4), np.float32)
    pos[:, :3] = (rng.random((N, 3)) - 0.5) * L
    pos[:, :3] = np.clip(pos[:, :3], -L / 2, np.nextafter(np.float32(L / 2), np.float32(0)))
    types = rng.integers(0, ntypes, size=N).astype(np.int32)
    pos[:, 3] = types.view(np.float32)
    return pos


KA_PARAMS = dict(  # Kob-Andersen 80:20 (BASELINE config 3)
    epsilon=[[1.0, 1.5], [1.5, 0.5]], sigma=[[1.0, 0.8], [0.8, 0.88]], r_cut=[[2.5, 2.0], [2.0, 2.2]])


KG_PARAMS = dict(  # Kremer-Grest bead-spring melt (BASELINE config 4)
    rho=0.85, r_cut=2.0 ** (1.0 / 6.0), fene_k=30.0, fene_r0=1.5, epsilon=1.0, sigma=1.0, kT=1.0)


def polymer_lattice(n_side, chain_len, rho=0.85, kT=1.0, seed=42):
    """n_side^3 beads on a simple-cubic lattice (a = rho^(-1/3)), threaded into linear chains of chain_len beads.

    The lattice sites are visited along a boustrophedon (snake) path so that consecutive sites are always lattice
    neighbors (bond length a = 1.056 at rho = 0.85 < FENE r0 = 1.5); the path is cut every chain_len sites.  With
    n_side = 100 and chain_len = 100 every chain is one straight row (10,000 chains x 100 beads = 1M beads).
    Returns (pos4, vel4, L, bonds[nbonds, 2] of particle tags).
    """
    N = n_side ** 3
    assert N % chain_len == 0, "chain_len must divide n_side^3"
    a = (1.0 / rho) ** (1.0 / 3.0)
    pos, vel, L = sc_lattice(n_side, a, kT=kT, seed=seed)
    # site index along the snake path: z slowest, y snakes within z, x snakes within y
    z, y, x = np.meshgrid(np.arange(n_side), np.arange(n_side), np.arange(n_side), indexing="ij")
    y = np.where(z % 2 == 1, n_side - 1 - y, y)
    row = z * n_side + np.where(z % 2 == 1, n_side - 1 - y, y)  # rows in visiting order
    x = np.where(row % 2 == 1, n_side - 1 - x, x)
    # sc_lattice numbers sites as (ix * n + iy) * n + iz (meshgrid 'ij'): take the snake's sites in path order
    site = ((x * n_side + y) * n_side + z).ravel()
    pos = np.ascontiguousarray(pos[site])
    p = np.arange(N, dtype=np.uint32)
    keep = (p % chain_len) != (chain_len - 1)
    bonds = np.stack([p[keep], p[keep] + 1], axis=1).astype(np.uint32)
    return pos, vel, L, bonds
