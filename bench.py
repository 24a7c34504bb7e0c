#!/usr/bin/env python
"""Benchmark of the MD hot path (BASELINE.json metric: timesteps/sec for LJ liquid).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload lj64k|lj1m|ka256k|lj8m]

One bench "step" = one System::run of --md-steps MD timesteps (NVE, LJ r_cut 3.0, r_buff 0.4, dt 0.005) of the synthetic
LJ liquid (N = 64,000, rho = 0.2).  BOTH arms start from the same equilibrated snapshot
tests/golden/lj64k_equil.npz (made by the unmodified reference: lattice start + 10,000 Langevin steps,
tests/golden/make_lj64k_equil.py).
`value` = MD timesteps / second with the state resident in HBM (CUDA events on the engine stream);
`e2e` = the same through the public host-buffer API (hoomd_b200.System: set_particles H2D, run, particles()+thermo D2H).
N > 1 (torchrun): BASELINE configs[4], LJ liquid 8M particles domain-decomposed over the N GPUs; before timing every rank
takes part in a 64k parity pass against the reference's CPU classes (tests/mg_check.py:parity_pass) whose result rides in
the line (parity_checked, max_force_err); rank 0 also measures the same 8M system on ONE GPU with the same protocol, so
`strong_scaling.efficiency` = tps_N / (N * tps_1) is reproducible from the record.
--impl reference times the reference's own CPU classes (oracle/_ref, HOOMD-blue 2.3.0 SINGLE_PRECISION, serial: the MD
CPU path has no threading and there is no MPI on the box) on a bounded sample of the same workload; it never loads
libhoomd_b200.so.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (n_side of the sc lattice, rho, kT, description)
    "lj64k": (40, 0.2, 1.2, "LJ liquid 64,000 particles, rho 0.2, r_cut 3.0, r_buff 0.4, NVE dt 0.005"),
    "lj1m": (100, 0.2, 1.2, "LJ liquid 1,000,000 particles, rho 0.2, r_cut 3.0, r_buff 0.4, NVE dt 0.005"),
    "lj8m": (200, 0.2, 1.2, "LJ liquid 8,000,000 particles, rho 0.2, r_cut 3.0, r_buff 0.4, NVE dt 0.005"),
    # BASELINE configs[3]: 10,000 chains x 100 beads, WCA (LJ r_cut 2^(1/6), shift) + FENE (K 30, r0 1.5), bond exclusions
    "kg1m": (100, 0.85, 1.0, "Kremer-Grest polymer melt 1,000,000 beads (10,000 x 100), rho 0.85, WCA + FENE, r_buff 0.4, "
                             "NVE dt 0.005"),
    # BASELINE configs[2]: 80:20 binary mixture, 256,000 particles (fcc 40^3 start), per-pair r_cut, Langevin
    "ka256k": (40, 1.2, 1.0, "Kob-Andersen 80:20 binary LJ 256,000 particles, rho 1.2, r_cut 2.5 sigma_ab (shifted), r_buff 0.4, "
                             "Langevin kT 1.0 dt 0.005"),
}
EQUIL_SNAPSHOT = os.path.join(ROOT, "tests", "golden", "lj64k_equil.npz")


def workloads_module():
    """hoomd_b200/workloads.py (pure numpy) loaded by path: the reference arm must not import the hoomd_b200 package"""
    import importlib.util
    spec = importlib.util.spec_from_file_location("_b200_workloads", os.path.join(ROOT, "hoomd_b200", "workloads.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def n_particles(workload):
    n_side = WORKLOADS[workload][0]
    return 4 * n_side ** 3 if workload.startswith("ka") else n_side ** 3


def default_md_steps(N):
    return 2000 if N <= 100000 else (400 if N <= 2000000 else 200)


def make_config(workload, world, grid=None):
    """the `config` both arms print (what is measured, not how one arm runs it)"""
    N = n_particles(workload)
    if world > 1 and grid is None:
        L = (N / WORKLOADS[workload][1]) ** (1.0 / 3.0)
        # DomainDecomposition::findDecomposition for a cube (hoomd/DomainDecomposition.cc:323-372): pure arithmetic
        best, area = None, None
        for nx in range(1, world + 1):
            for ny in range(1, world // nx + 1):
                nz = world // (nx * ny)
                if nx * ny * nz != world:
                    continue
                a = L * L * ((nz - 1) + (ny - 1) + (nx - 1))
                if area is None or a < area:
                    best, area = (nx, ny, nz), a
        grid = best
    return {"workload": WORKLOADS[workload][3], "N": N,
            "parallelism": "1 domain" if world == 1 else
            "spatial domain decomposition %s over %d GPUs, ghosts + migration over NVLink peer memory" % (
                "x".join(str(g) for g in grid), world),
            "start_state": ("equilibrated snapshot tests/golden/lj64k_equil.npz (made by the reference: sc lattice + 10,000 "
                            "Langevin steps kT 1.2)") if workload == "lj64k" else
            "lattice start, Gaussian velocities (numpy default_rng(42)); the CUDA arm melts it first (untimed)",
            "l2": "L2 flushed (256 MiB write) between timed iterations" + (
                "; within an iteration the 64k working set is L2-resident, which is the workload"
                if N <= 100000 else "; per-GPU working set exceeds L2"),
            "nlist": {"r_buff": 0.4, "check_period": 1}}
KG_CHAIN = 100


METHOD_OVERRIDE = os.environ.get("HOOMD_B200_BENCH_METHOD", "")


def n_types(workload):
    return 2 if workload.startswith("ka") else 1


def add_method(s, workload, kT):
    """the integration method the BASELINE config names: Langevin for Kob-Andersen, NVE otherwise"""
    if METHOD_OVERRIDE == "nvt":  # A/B measurement of the thermostat's cost, never the default line
        s.add_nvt(kT, 0.5)
    elif workload.startswith("ka"):
        s.add_langevin(kT, 7)
    else:
        s.add_nve()
BYTES_PER_PARTICLE_STEP = 420.0   # SURVEY 8(d) planning figure (NVE, fused net force, amortised rebuild)
FLOPS_PER_PARTICLE_STEP = 1800.0
PAIR_BYTES_PER_PARTICLE = 172.0   # pair kernel: pos_i 16 + n_neigh 4 + head 4 + nlist 4*32.9 + force 16
PAIR_FLOPS_PER_PARTICLE = 1316.0  # 40 flop x 32.9 neighbors


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _loop(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([x.strip() for x in out.strip().split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._t = threading.Thread(target=self._loop, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        sm, smax, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                smax.append(float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


EQUIL_STEPS = None  # --equil-steps override (profiling runs)
USE_GRAPH = int(os.environ.get("HOOMD_B200_USE_GRAPH", "2"))  # 0 host-driven, 1 plain graph, 2 conditional graph
PAIR_TPP = int(os.environ.get("HOOMD_B200_PAIR_TPP", "0"))     # A/B sweeps only: lanes per particle (0 = the engine's choice)


def _lj_snapshot(workload):
    n_side, rho, kT, _ = WORKLOADS[workload]
    return workloads_module().sc_lattice(n_side, (1.0 / rho) ** (1.0 / 3.0), kT=kT, seed=42)


def _snapshot(workload):
    """(pos4, vel4, L, bonds or None)"""
    workloads = workloads_module()
    n_side, rho, kT, _ = WORKLOADS[workload]
    if workload.startswith("kg"):
        return workloads.polymer_lattice(n_side, KG_CHAIN, rho=rho, kT=kT, seed=42)
    if workload.startswith("ka"):
        pos, vel, L = workloads.fcc_lattice(n_side, rho, kT=kT, seed=42, ntypes=2, type_fractions=[0.8, 0.2])
        return pos, vel, L, None
    return _lj_snapshot(workload) + (None,)


def load_equil_snapshot():
    """the 64k start state of both arms (tag order)"""
    g = np.load(EQUIL_SNAPSHOT)
    return dict(pos=np.ascontiguousarray(g["pos"]), vel=np.ascontiguousarray(g["vel"]),
                image=np.ascontiguousarray(g["image"].astype(np.int32)), bonds=None), float(g["L"])


def setup_forces(s, workload, bonds, **kw):
    """md.nlist.cell + the workload's force field on any implementation (hoomd_b200.System, RefSystem, OracleSystem)"""
    if workload.startswith("kg"):
        s.set_bonds(bonds)
        s.setup_lj(r_cut=2.0 ** (1.0 / 6.0), r_buff=0.4, shift_mode=1, **kw)
        s.setup_fene(k=30.0, r0=1.5)
        s.exclusions_from_bonds(True)
    elif workload.startswith("ka"):
        s.setup_lj(shift_mode=1, r_buff=0.4, **workloads_module().KA_PARAMS, **kw)
    else:
        s.setup_lj(r_cut=3.0, r_buff=0.4, **kw)


def make_state(workload, rank=0):
    """lj64k: the committed equilibrated snapshot.  Other workloads: lattice start + Langevin equilibration ON THE GPU
    ENGINE (not timed); returns host arrays + box length."""
    if workload == "lj64k" and EQUIL_STEPS is None:
        return load_equil_snapshot()
    import hoomd_b200
    _, _, kT, _ = WORKLOADS[workload]
    pos, vel, L, bonds = _snapshot(workload)
    s = hoomd_b200.System(pos, (L, L, L), vel4=vel, ntypes=n_types(workload))
    setup_forces(s, workload, bonds)
    nequil = EQUIL_STEPS if EQUIL_STEPS is not None else (10000 if len(pos) <= 100000 else 2000)
    s.set_sorter(300)
    s.set_tuning(use_graph=USE_GRAPH)
    if workload.startswith("ka"):
        # the compressed sc start (rho 1.2) relaxes with a short time step first
        s.setup_integrator(0.002)
        s.add_langevin(kT, 42)
        s.run(nequil)
    s.setup_integrator(0.005)
    s.add_langevin(kT, 42)
    s.run(nequil)
    p = s.particles()
    s.close()
    # back to TAG order: the bond table names particles by tag (= index of a freshly created system)
    order = np.argsort(p["tag"], kind="stable")
    p = {k: np.ascontiguousarray(v[order]) for k, v in p.items()}
    p["bonds"] = bonds
    return p, L


def new_system(state, L, device=0, workload="lj64k"):
    import hoomd_b200
    s = hoomd_b200.System(state["pos"], (L, L, L), vel4=state["vel"], image=state["image"], device=device,
                          ntypes=n_types(workload))
    setup_forces(s, workload, state.get("bonds"))
    s.setup_integrator(0.005)
    add_method(s, workload, WORKLOADS[workload][2])
    s.set_sorter(300)  # hoomd.update.sort default period on the GPU (hoomd/update.py:258-263)
    s.set_tuning(pair_tpp=PAIR_TPP, use_graph=USE_GRAPH)
    return s


def new_domain_system(workload, local_rank):
    """N > 1: the same lattice snapshot on every rank, each keeps its sub-domain; Langevin equilibration (untimed,
    tag-keyed noise => independent of the decomposition), then the NVE integrator on the same device-resident state."""
    from hoomd_b200 import domain
    _, _, kT, _ = WORKLOADS[workload]
    pos, vel, L = _lj_snapshot(workload)
    s = domain.DomainSystem(pos, (L, L, L), r_ghost=3.4, vel4=vel, device=local_rank)
    del pos, vel
    s.setup_lj(r_cut=3.0, r_buff=0.4)
    s.setup_integrator(0.005)
    s.add_langevin(kT, 42)
    s.set_sorter(300)
    s.set_tuning(use_graph=USE_GRAPH)
    s.run(EQUIL_STEPS if EQUIL_STEPS is not None else 2000)
    s.add_nve()
    return s, L


def run_reference(args):
    """reference arm: the reference's own CPU classes on the box's host cores, bounded sample per step.  Never imports the
    hoomd_b200 package / loads libhoomd_b200.so: the 64k start state is the committed snapshot, other workloads start from
    the lattice."""
    from oracle import ref, oracle_c
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    workload = args.workload or ("lj64k" if world == 1 else "lj8m")  # the same choice as the CUDA arm
    n_side, rho, kT, desc = WORKLOADS[workload]
    N = n_particles(workload)
    md_steps = args.ref_md_steps or (20 if N <= 100000 else (2 if N <= 2000000 else 1))
    if workload == "lj64k":
        state, L = load_equil_snapshot()
    else:
        pos, vel, L, bonds = _snapshot(workload)
        state = dict(pos=pos, vel=vel, image=np.zeros((len(pos), 3), np.int32), bonds=bonds)
    nbt = 1 if state.get("bonds") is not None else 0
    if ref.available():
        s = ref.RefSystem(state["pos"], (L, L, L), vel4=state["vel"], image=state["image"], n_bond_types=nbt,
                          ntypes=n_types(workload))
        setup_forces(s, workload, state.get("bonds"), full=False)  # CPU md.pair.lj default: half list + Newton's 3rd law
        kind, cores = "reference", 1
        s.set_sorter(100)  # hoomd.update.sort default period on the CPU
    else:
        cores = os.cpu_count() or 1
        s = oracle_c.OracleSystem(state["pos"], (L, L, L), vel4=state["vel"], image=state["image"], threads=cores,
                                  n_bond_types=nbt, ntypes=n_types(workload))
        setup_forces(s, workload, state.get("bonds"))
        kind = "port"
    s.setup_integrator(0.005)
    add_method(s, workload, WORKLOADS[workload][2])
    for _ in range(args.warmup):
        s.run(md_steps)
    t = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        s.run(md_steps)
        t.append(time.perf_counter() - t0)
    total = sum(t)
    tps = md_steps * args.steps / total
    sample = "%d MD steps per bench step x %d steps of the %d-particle workload, %s" % (
        md_steps, args.steps, N, "HOOMD-blue 2.3.0 CPU classes (oracle/_ref), half neighbor list, 1 core (no MPI on the box)"
        if kind == "reference" else "C restatement (oracle/oracle.c) with OpenMP")
    line = {"impl": "reference", "metric": "timesteps/sec", "value": tps, "unit": "timesteps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": make_config(workload, world),
            "md_steps_per_step": md_steps,
            "particle_steps_per_sec": tps * N,
            "cpu_baseline": {"value": tps, "unit": "timesteps/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": tps, "unit": "timesteps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def cpu_baseline(state, L, budget_s=15.0, workload="lj64k"):
    """bounded sample of the same workload on the host cores with the reference's own CPU classes"""
    from oracle import ref, oracle_c
    N = len(state["pos"])
    nbt = 1 if state.get("bonds") is not None else 0
    if ref.available():
        s = ref.RefSystem(state["pos"], (L, L, L), vel4=state["vel"], image=state["image"], n_bond_types=nbt,
                          ntypes=n_types(workload))
        setup_forces(s, workload, state.get("bonds"), full=False)
        kind, cores = "reference", 1
    elif oracle_c.available():
        cores = os.cpu_count() or 1
        s = oracle_c.OracleSystem(state["pos"], (L, L, L), vel4=state["vel"], image=state["image"], threads=cores,
                                  n_bond_types=nbt, ntypes=n_types(workload))
        setup_forces(s, workload, state.get("bonds"))
        kind = "port"
    else:
        return None
    s.setup_integrator(0.005)
    add_method(s, workload, WORKLOADS[workload][2])
    t0 = time.perf_counter()
    s.run(5)
    per = (time.perf_counter() - t0) / 5
    n = max(5, min(2000, int(budget_s / max(per, 1e-6))))
    t0 = time.perf_counter()
    s.run(n)
    dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": "timesteps/s", "cores": cores, "kind": kind,
            "sample": "%d MD steps of the same %d-particle state (%.1f s), %s" % (
                n, N, dt, "HOOMD-blue 2.3.0 CPU classes built from /root/reference (oracle/_ref), half list, serial"
                if kind == "reference" else "C restatement oracle/oracle.c, OpenMP")}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS),
                    help="default: lj64k on 1 GPU (BASELINE configs[1]), lj8m domain-decomposed on N > 1 (configs[4])")
    ap.add_argument("--no-scale-ref", action="store_true", help="skip the 1-GPU 8M-particle reference point")
    ap.add_argument("--md-steps", type=int, default=0, help="MD timesteps per bench step (0 = by workload)")
    ap.add_argument("--ref-md-steps", type=int, default=0, help="MD timesteps per bench step of the reference arm (0 = by size)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--equil-steps", type=int, default=None)
    ap.add_argument("--no-flush", action="store_true", help="profiling only: skip the L2 flush between timed iterations")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the parity pass against the reference classes")
    ap.add_argument("--parity-steps", type=int, default=100, help="NVE steps of the N > 1 parity pass")
    args = ap.parse_args()
    global EQUIL_STEPS
    EQUIL_STEPS = args.equil_steps

    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- hoomd_b200 has no CPU fallback")
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    workload = args.workload or ("lj64k" if world == 1 else "lj8m")
    n_side, rho, kT, desc = WORKLOADS[workload]
    N = n_particles(workload)
    md_steps = args.md_steps or default_md_steps(N)
    hbm_peak, peak_src = peaks()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    def barrier(collective=True):
        torch.cuda.synchronize()
        if world > 1 and collective:
            dist.barrier()
        torch.cuda.synchronize()

    def measure(sys_, steps_per_run, collective=True):
        """W untimed warm-up runs, then K timed runs bracketed by barrier + synchronize; CUDA-event milliseconds per run on
        the engine stream, summed and (domain-decomposed) taken as the max over ranks"""
        for _ in range(args.warmup):
            sys_.run(steps_per_run)
        barrier(collective)
        ms_runs, n_launch = [], 0
        t0 = time.perf_counter()
        for _ in range(args.steps):
            if not args.no_flush:
                flush.fill_(1)       # evict L2 between timed iterations (the 15 MB working set re-warms in < 1 MD step)
                torch.cuda.synchronize()
            _, ms = sys_.run(steps_per_run)
            ms_runs.append(ms)
            n_launch += sys_.stats()["kernels_last_run"]
        barrier(collective)
        wall = time.perf_counter() - t0
        tot = sum(ms_runs)
        if world > 1 and collective:
            t = torch.tensor([tot], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            tot = float(t.item())
        return tot, ms_runs, n_launch, wall

    def measure_e2e(sys_, steps_per_run, collective=True):
        """the same through the host-buffer API: H2D of the step's inputs, run, D2H of the result, every bench step"""
        host = sys_.particles()
        e2e_t, h2d, d2h, th = [], 0, 0, None
        # one untimed pass first, like the device-timed arm's warm-up (the first pass through the host-buffer API pays
        # one-time costs -- staging buffers, graph re-instantiation: 65-110 ms against 57 ms at 64k)
        for it in range(-1, max(1, args.steps)):
            t0 = time.perf_counter()
            sys_.set_particles(host["pos"], host["vel"], host["image"])     # H2D of the step's inputs
            sys_.run(steps_per_run)
            host = sys_.particles()                                          # D2H of the step's result
            th = sys_.thermo()
            if it >= 0: e2e_t.append(time.perf_counter() - t0)
            h2d = host["pos"].nbytes + host["vel"].nbytes + host["image"].nbytes
            d2h = sum(host[k].nbytes for k in host) + 5 * 8
        tot = sum(e2e_t)
        if world > 1 and collective:
            t = torch.tensor([tot], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            tot = float(t.item())
        return {"value": steps_per_run * len(e2e_t) / tot, "unit": "timesteps/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "thermo_T": th["T"], "ms_runs": [1e3 * x for x in e2e_t]}

    # ---- N > 1: parity of the domain-decomposed path against the reference's CPU classes, before anything is timed -------
    parity = None
    if world > 1 and not args.no_parity:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import mg_check
        st64, L64 = load_equil_snapshot()
        parity = mg_check.parity_pass(st64["pos"], st64["vel"], st64["image"], L64, args.parity_steps, local_rank,
                                      check_ghosts=False, verbose=False)
        parity.pop("mg_stats", None)
        parity.pop("mg_timing", None)
        parity["workload"] = ("LJ liquid 64,000 (tests/golden/lj64k_equil.npz) on the same %d-GPU decomposition vs HOOMD-blue "
                              "2.3.0 CPU classes: forces / energies within 1e-5 of sum|pair terms|, neighbor-list row lengths "
                              "equal, %d NVE steps with equal rebuild counts" % (world, args.parity_steps))

    if world == 1:
        state, L = make_state(workload, rank)
        s = new_system(state, L, device=local_rank, workload=workload)
    else:
        state = None
        s, L = new_domain_system(workload, local_rank)

    # ---- device-resident throughput ---------------------------------------------------------------------------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    total_ms, gpu_ms, launches, t_wall = measure(s, md_steps)
    clocks = sampler.stop()
    tps_rank = md_steps * args.steps / (total_ms * 1e-3)
    tps = tps_rank  # one domain-decomposed system: every rank advances the same timesteps
    stats = s.stats()
    n_local = s.N

    # ---- dominant kernel: the single-kernel MD step (pair force + both half steps + distance check), timed live with
    # CUDA events on the engine stream (b200_sys_time_kernel); algorithmic bytes per particle per launch:
    # read pos 16 + n_neigh 4 + head 4 + nlist 4*nbar + vel 16 + image 12 + last_pos 16; write pos 16 + vel 16 + accel 12
    # + image 12  = 124 + 4*nbar
    s.time_kernel(2, 2)  # first launches of a kernel pay for lazy module loading
    step_ms = s.time_kernel(2, 20)
    one_step_ms, multistep = step_ms, bool(world == 1 and s.stepping()["multistep"])
    if multistep:
        # the run went through k_md_steps (several MD steps per launch): its per-step time incl. the grid-wide barrier
        step_ms = s.time_kernel(5, 40)
    pair_ms = s.time_kernel(0, 200)
    nlist_ms = s.time_kernel(1, 20)
    cell_only_ms, nlist_only_ms = s.time_kernel(3, 20), s.time_kernel(4, 20)
    n_neigh = s.nlist(counts_only=True)[0][:n_local]
    nbar = float(n_neigh.mean())
    step_bytes = (124.0 + 4.0 * nbar) * n_local
    if workload.startswith("kg"):
        # + k_bond_forces (pos 16 + tag 4 + n_bonds 4 + 2 bonds x (entry 8 + rtag 4) + force 16 = 64) and the 16 B read of
        # that force by k_md_step; time_kernel(2) launches both per step
        step_bytes += 80.0 * n_local
    step_flops = (40.0 * nbar + 55.0) * n_local
    achieved = step_bytes / (step_ms * 1e-3) / 1e9
    # DRAM bytes per launch from the committed ncu --set full capture of this workload (single GPU only), else null
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            ent = json.load(f).get(workload) if world == 1 else None
        if ent:
            traffic, traffic_src = ent["bytes"], ent["source"]
    except (OSError, ValueError):
        pass
    roofline = {"bound": "hbm", "kernel": ("k_md_steps<EvaluatorPairLJ> (several MD steps per launch; per step: force + step "
                                          "two + step one + distance check + grid barrier)" if multistep else
                                          "k_md_step<EvaluatorPairLJ> (force + step two + step one + distance check)")
                + (" + k_bond_forces<EvaluatorBondFENE>" if workload.startswith("kg") else ""),
                "one_step_kernel_ms": one_step_ms,
                "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": traffic,
                "traffic_source": traffic_src,
                "peak_source": peak_src, "kernel_ms": step_ms, "algorithmic_bytes_per_launch": step_bytes,
                "mean_neighbors": nbar,
                "fp32_tflops": step_flops / (step_ms * 1e-3) / 1e12,
                "fp32_peak_tflops": 148 * 128 * 2 * 1.965e9 / 1e12,
                "pair_kernel_ms": pair_ms, "pair_kernel_gbs": (40.0 + 4.0 * nbar) * n_local / (pair_ms * 1e-3) / 1e9,
                "nlist_rebuild_ms": nlist_ms, "cell_list_ms": cell_only_ms, "nlist_kernel_ms": nlist_only_ms,
                "step_hbm_frac": BYTES_PER_PARTICLE_STEP * n_local * tps_rank / (hbm_peak * 1e9),
                "note": ("working set at 64k is L2-resident (15 MB < 126 MB L2): the step is launch/latency bound, "
                         "the HBM fraction is low by construction (SURVEY 8d)") if N <= 100000 else
                        "per-rank figures (local particles of rank 0)"}

    # ---- end to end through the host-buffer API ----------------------------------------------------------------------
    e2e = measure_e2e(s, md_steps)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(state, L, workload=workload)

    mg_stats = None
    if world > 1:
        mg_stats = s.mg_stats()
        mg_stats["grid"] = list(s.grid)
        mg_stats["barrier_wait"] = s.mg_timing()  # device-measured: arrival at a barrier -> every rank seen
    s.close()
    s = None

    # ---- the 1-GPU point of the multi-GPU workload (BASELINE configs[4]: parallel efficiency = TPS_n / (n * TPS_1) with
    # TPS_1 measured at 8M on one GPU), SAME protocol as the N > 1 runs (warm-up W, K x md_steps, sum of the
    # CUDA-event times, e2e).  At N = 1 it rides beside the 64k headline as `scale_ref`; at N > 1 rank 0 measures it on its
    # own GPU after the timed region (the other ranks wait), so every line carries its own denominator.
    scale_ref, strong = None, None
    want_ref = (world == 1 and workload == "lj64k") or (world > 1 and workload == "lj8m")
    if want_ref and not args.no_scale_ref:
        if rank == 0:
            st8, L8 = make_state("lj8m")
            s8 = new_system(st8, L8, device=local_rank, workload="lj8m")
            md8 = default_md_steps(n_particles("lj8m"))
            tot8, ms8, _, _ = measure(s8, md8, collective=False)
            tps8 = md8 * args.steps / (tot8 * 1e-3)
            scale_ref = {"workload": WORKLOADS["lj8m"][3], "n_gpus": 1, "value": tps8, "unit": "timesteps/s",
                         "particle_steps_per_sec": tps8 * 8e6, "md_steps_per_step": md8, "steps": args.steps,
                         "warmup": args.warmup, "ms_runs": ms8, "k_md_step_ms": s8.time_kernel(2, 10),
                         "nlist_rebuild_ms": s8.time_kernel(1, 5),
                         "e2e": measure_e2e(s8, md8, collective=False)}
            s8.close()
            if world > 1:
                strong = {"workload": "lj8m", "tps_1": tps8, "tps_n": tps, "n": world, "efficiency": tps / (world * tps8),
                          "e2e_efficiency": e2e["value"] / (world * scale_ref["e2e"]["value"]),
                          "how": "tps_1: the same 8M-particle system on ONE GPU (rank 0's, after the timed region), same "
                                 "warm-up / steps / md_steps protocol; efficiency = tps_n / (n * tps_1)"}
        if world > 1:
            dist.barrier()

    if rank == 0:
        cfg = make_config(workload, world, mg_stats["grid"] if mg_stats else None)
        if METHOD_OVERRIDE:
            cfg["workload"] += " [integration method overridden: %s]" % METHOD_OVERRIDE
        line = {"metric": "timesteps/sec", "value": tps, "unit": "timesteps/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": cfg,
                "md_steps_per_step": md_steps,
                "engine": {"sorter_period": 300,
                           "stepping": {0: "host-driven", 1: "CUDA graph, early-exit rebuild kernels",
                                        2: "CUDA graph, conditional IF node",
                                        3: "stream launches of the graph kernels"}[USE_GRAPH],
                           "multistep_kernel": multistep,
                           "rebuilds_total": stats["updates"] + stats["forced"], "Nmax": stats["Nmax0"]},
                "particle_steps_per_sec": tps * N,
                "wall_tps": md_steps * args.steps / t_wall,
                "gpu_launches": launches, "ms_runs": gpu_ms, "nlist_overflows": stats["overflows"],
                "clocks": clocks,
                "roofline": roofline,
                "e2e": e2e,
                "cpu_baseline": cpu}
        if scale_ref:
            line["scale_ref"] = scale_ref
        if strong:
            line["strong_scaling"] = strong
        if parity:
            line["parity_checked"] = True
            line["max_force_err"] = parity["max_force_err"]
            line["parity"] = parity
        if mg_stats:
            line["domain"] = mg_stats
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
