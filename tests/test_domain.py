"""Host side of the domain decomposition (hoomd_b200/domain.py): CPU tests, including a world_size-2 gloo run.
Reference behaviour: hoomd/DomainDecomposition.cc:323-372 (findDecomposition), :531-557 (calculateLocalBox),
:566-613 (placeParticle)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _domain():
    # import without touching the CUDA library
    from hoomd_b200 import domain
    return domain


def test_find_decomposition_matches_reference_choices():
    d = _domain()
    # a cube: 2 -> 1x1x2, 4 -> 1x2x2, 8 -> 2x2x2 (the loop order of DomainDecomposition.cc:343-366 keeps the first minimum)
    assert d.find_decomposition((10, 10, 10), 1) == (1, 1, 1)
    assert d.find_decomposition((10, 10, 10), 2) == (1, 1, 2)
    assert d.find_decomposition((10, 10, 10), 4) == (1, 2, 2)
    assert d.find_decomposition((10, 10, 10), 8) == (2, 2, 2)
    # a slab: cut the long axis
    assert d.find_decomposition((40, 10, 10), 4) == (4, 1, 1)
    assert d.find_decomposition((10, 40, 10), 2) == (1, 2, 1)
    for n in (1, 2, 3, 4, 6, 8):
        g = d.find_decomposition((7, 9, 11), n)
        assert g[0] * g[1] * g[2] == n


def test_place_particles_partitions_the_box():
    d = _domain()
    rng = np.random.default_rng(0)
    L = (12.0, 14.0, 16.0)
    pos = ((rng.random((5000, 3)) - 0.5) * np.array(L)).astype(np.float32)
    for grid in [(1, 1, 2), (1, 2, 2), (2, 2, 2), (4, 1, 1)]:
        owner = d.place_particles(pos, L, grid)
        world = grid[0] * grid[1] * grid[2]
        assert owner.min() >= 0 and owner.max() < world
        total = 0
        for r in range(world):
            lo, hi, per = d.local_bounds(L, grid, r)
            mine = pos[owner == r]
            total += len(mine)
            for a in range(3):
                assert (mine[:, a] >= lo[a]).all() and (mine[:, a] < hi[a]).all()
                assert per[a] == (1 if grid[a] == 1 else 0)
        assert total == len(pos)
    # a particle exactly on a cut plane belongs to the upper domain (lo <= x < hi)
    p = np.array([[0.0, 0.0, 0.0]], np.float32)
    assert d.place_particles(p, L, (2, 2, 2))[0] == 7


def test_rank_grid_mapping_roundtrip():
    d = _domain()
    grid = (2, 3, 4)
    seen = set()
    for r in range(24):
        ix, iy, iz = d.rank_to_grid(r, grid)
        assert (iz * grid[1] + iy) * grid[0] + ix == r
        seen.add((ix, iy, iz))
    assert len(seen) == 24


_WORKER = r'''
import os, sys
import numpy as np
import torch.distributed as dist
sys.path.insert(0, %(root)r)
from hoomd_b200 import domain, workloads
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
pos, vel, L = workloads.sc_lattice(10, 1.7, kT=1.0, seed=3)
grid = domain.find_decomposition((L, L, L), world)
owner = domain.place_particles(pos[:, :3], (L, L, L), grid)
mine = np.nonzero(owner == rank)[0]
counts = [None] * world
dist.all_gather_object(counts, int(len(mine)))
assert sum(counts) == len(pos), counts
tags = [None] * world
dist.all_gather_object(tags, mine)
allt = np.sort(np.concatenate(tags))
assert np.array_equal(allt, np.arange(len(pos)))
lo, hi, per = domain.local_bounds((L, L, L), grid, rank)
bounds = [None] * world
dist.all_gather_object(bounds, (lo, hi))
# the domains tile the box: along the split axis hi of one is lo of the next
for a in range(3):
    los = sorted(set(float(b[0][a]) for b in bounds)); his = sorted(set(float(b[1][a]) for b in bounds))
    assert abs(los[0] + L / 2) < 1e-5 and abs(his[-1] - L / 2) < 1e-5
    assert los[1:] == his[:-1]
dist.barrier()
if rank == 0:
    print("DOMAIN_GLOO_OK", grid, counts)
dist.destroy_process_group()
'''


def test_world_size_2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER % {"root": ROOT})
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                         capture_output=True, text=True, timeout=300)
    assert "DOMAIN_GLOO_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


@pytest.mark.gpu
def test_multigpu_parity_vs_reference():
    """tests/mg_check.py under torchrun on 2 GPUs: ghosts, forces, list row lengths, migration, rebuild counts and
    trajectories of the domain-decomposed engine vs the reference's CPU classes (oracle/_ref)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", str(port),
                          os.path.join(ROOT, "tests", "mg_check.py"), "24", "300"],
                         capture_output=True, text=True, timeout=600)
    assert "MG_CHECK_OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]
