// ENGINE, cell-tile path (tile.cu): geometry selection, rebuild, host-side overflow handling.  Included by engine.cu.
//
// In this mode the NeighborList object keeps its whole state machine (needsUpdating, distance check, last_pos, the
// counters) but stores the list as 16-bit slots into per-CTA shared-memory tiles, and every rebuild re-orders the
// particle data by cell.  Single GPU, orthorhombic box, >= 3 cells per periodic axis; otherwise the engine stays on the
// CSR walk (pair_mode 0).
namespace b200
{
static const unsigned int TILE_CAP_LIMIT = 6144; // staged positions per CTA: 96 kB of shared memory

static unsigned int round_up_u(unsigned int v, unsigned int m) { return (v + m - 1) / m * m; }

// pick cells-per-tile and the staging capacity from the mean cell population
static bool tile_pick_geometry(b200_system* s)
    {
    if (s->mg) return false;
    if (s->dim.x < 3 || s->dim.y < 3 || s->dim.z < 3) return false; // 27 distinct adjacent cells needed
    for (int d = 0; d < 3; d++)
        if (!s->box.periodic[d]) return false;
    const double ncells = (double)s->dim.x * s->dim.y * s->dim.z;
    const double per_cell = (double)s->N / ncells;
    static const unsigned int shapes[][3] = {{4, 2, 2}, {2, 2, 2}, {2, 2, 1}, {2, 1, 1}, {1, 1, 1}};
    unsigned int first = 0;
    if (s->tile_req[0])
        {
        // explicit request (tuning): take it if it fits the kernel's static limits
        const unsigned int* r = s->tile_req;
        if ((r[0] + 2) * (r[1] + 2) * (r[2] + 2) > 128 || r[1] * r[2] > 16) return false;
        const double halo = (r[0] + 2.0) * (r[1] + 2.0) * (r[2] + 2.0) * per_cell;
        const unsigned int cap = round_up_u((unsigned int)(halo * 1.35 + 160.0), 64);
        if (cap > TILE_CAP_LIMIT) return false;
        s->tile_geom.b = make_uint3(r[0], r[1], r[2]);
        s->tile_geom.cap = cap;
        first = 99;
        }
    for (unsigned int k = first; k < 5; k++)
        {
        const unsigned int* r = shapes[k];
        const double halo = (r[0] + 2.0) * (r[1] + 2.0) * (r[2] + 2.0) * per_cell;
        const unsigned int cap = round_up_u((unsigned int)(halo * 1.35 + 160.0), 64);
        if (cap > 2560 && k < 4) continue; // keep several CTAs per SM resident; the smallest tile may use more
        if (cap > TILE_CAP_LIMIT) return false;
        s->tile_geom.b = make_uint3(r[0], r[1], r[2]);
        s->tile_geom.cap = cap;
        break;
        }
    s->tile_geom.dim = make_uint3(s->dim.x, s->dim.y, s->dim.z);
    const uint3 b = s->tile_geom.b;
    s->tile_geom.tdim = make_uint3((s->dim.x + b.x - 1) / b.x, (s->dim.y + b.y - 1) / b.y, (s->dim.z + b.z - 1) / b.z);
    return true;
    }

static void tile_alloc_list(b200_system* s)
    {
    s->tile_pitch = round_up_u(s->Ncap, 64);
    s->tile_Nmax = round_up_u(s->tile_Nmax, 8);
    s->tile_ell.alloc((size_t)(s->tile_Nmax / 8) * s->tile_pitch, false);
    }

// called when the pair potential / neighbor list parameters are known (setup_lj) and on tuning changes
static void tile_configure(b200_system* s)
    {
    s->tile_on = false;
    invalidate_graph(s);
    // auto (-1) resolves to the CSR walk: measured on the current tree the tile path loses at every size (LJ 1M: step kernel
    // 153 vs 133 us, rebuild 0.84 vs 0.59 ms; DESIGN.md section 4b, profiles/r2w_*) -- explicit opt-in only
    if (!s->have_nlist || s->pair_mode != 1) return;
    if (s->bond_kind >= 0 || s->exclusions_set)
        throw std::runtime_error("tile: the cell-tile path carries neither bonded forces nor nlist exclusions");
    if (s->pair_kind != B200_PAIR_LJ) throw std::runtime_error("tile: the cell-tile path is instantiated for md.pair.lj only");
    if (!tile_pick_geometry(s))
        {
        if (s->pair_mode == 1) throw std::runtime_error("tile: cell-tile path needs one GPU, a periodic orthorhombic box with "
                                                        ">= 3 cells per axis and <= 6144 particles per staged tile");
        return;
        }
    if (!s->tile_tpp) s->tile_tpp = s->N <= 200000 ? 2 : 1; // lanes per owned particle
    // expected neighbors within r_list at the mean density, with head room; grown on overflow like Nmax
    const double V = (double)s->box.L[0] * s->box.L[1] * s->box.L[2];
    const double rl = s->rcut_max_max + s->r_buff;
    const double n_est = 4.18879 * rl * rl * rl * (double)s->N / V;
    s->tile_Nmax = round_up_u((unsigned int)(n_est * 1.3 + 16.0), 8);
    if (const char* v = getenv("HOOMD_B200_TILE_NMAX0")) s->tile_Nmax = std::max(1, atoi(v)); // tests: force the growth path
    tile_alloc_list(s);
    const unsigned int C = s->Ncap;
    s->tile_nn.alloc(C);
    s->tile_perm.alloc(C);
    s->tile_cond.alloc(2);
    if (s->vel_alt.n < C) s->vel_alt.alloc(C);
    if (s->accel_alt.n < C) s->accel_alt.alloc(C);
    if (s->image_alt.n < C) s->image_alt.alloc(C);
    if (s->tag_alt.n < C) s->tag_alt.alloc(C);
    s->tile_on = true;
    s->force_update = true;
    }

static void tile_build_args(b200_system* s, TileBuildArgs& a, const unsigned int* enable, const float4* pos)
    {
    memset(&a, 0, sizeof(a));
    a.enable = enable;
    a.geom = s->tile_geom;
    a.cell_start = s->cell_start.p;
    a.pos = pos;
    a.ell4 = s->tile_ell.p;
    a.pitch = s->tile_pitch; a.Nmax = s->tile_Nmax;
    a.nn = s->tile_nn.p;
    a.conditions = s->tile_cond.p;
    a.r_cut = s->d_r_cut.p; a.r_listsq = s->d_r_listsq.p; a.ntypes = s->ntypes;
    a.box = make_box(s->box);
    }

static void tile_step_args(b200_system* s, TileStepArgs& a, int force_only)
    {
    memset(&a, 0, sizeof(a));
    a.geom = s->tile_geom;
    a.cell_start = s->cell_start.p;
    a.ell4 = s->tile_ell.p;
    a.pitch = s->tile_pitch; a.tpp = s->tile_tpp; a.Nmax = s->tile_Nmax;
    a.nn = s->tile_nn.p;
    a.force_only = force_only;
    }

// CellList (with the index permutation) -> re-order the particle data by cell -> tile-slot lists.  `pos` is the
// position buffer of the ping-pong phase; it holds the cell-sorted positions afterwards.
static void tile_enqueue_rebuild(b200_system* s, const unsigned int* enable, float4* pos)
    {
    b200_cell_list_args ca;
    cell_list_args(s, ca);
    ca.d_pos = (const float*)pos;
    CUDA_OK(launch_cell_list(&ca, enable, s->stream));
    TileSortArgs sa;
    memset(&sa, 0, sizeof(sa));
    sa.enable = enable; sa.N = s->N; sa.perm = s->tile_perm.p; sa.sorted_pos = s->cell_xyzf.p;
    sa.pos = pos; sa.last_pos = s->last_pos.p;
    sa.vel = s->vel.p; sa.vel_alt = s->vel_alt.p;
    sa.accel = (float*)s->accel.p; sa.accel_alt = (float*)s->accel_alt.p;
    sa.image = (int*)s->image.p; sa.image_alt = (int*)s->image_alt.p;
    sa.tag = s->tag.p; sa.tag_alt = s->tag_alt.p; sa.rtag = s->rtag.p;
    CUDA_OK(launch_tile_sort(sa, s->stream));
    TileBuildArgs ba;
    tile_build_args(s, ba, enable, pos);
    CUDA_OK(launch_tile_build(ba, s->stream));
    }

// NeighborList::checkConditions for the tile lists: grow the row capacity / the staging capacity on the host
static bool tile_check_conditions(b200_system* s)
    {
    unsigned int c[2];
    s->tile_cond.download(c, 2, s->stream);
    bool result = false;
    if (c[1])
        {
        // some tile's halo population exceeded the staging capacity: enlarge it (or fall to smaller tiles)
        unsigned int cap = round_up_u((unsigned int)(c[1] * 1.15f) + 64, 64);
        while (cap > TILE_CAP_LIMIT)
            {
            uint3& b = s->tile_geom.b;
            if (b.x > 1 && b.x >= b.y && b.x >= b.z) b.x /= 2;
            else if (b.y > 1 && b.y >= b.z) b.y /= 2;
            else if (b.z > 1) b.z /= 2;
            else throw std::runtime_error("tile: cell population too large for the cell-tile path; use pair_mode 0");
            const double per_cell = (double)s->N / ((double)s->dim.x * s->dim.y * s->dim.z);
            cap = round_up_u((unsigned int)((b.x + 2.0) * (b.y + 2.0) * (b.z + 2.0) * per_cell * 1.6 + 256.0), 64);
            }
        s->tile_geom.cap = cap;
        const uint3 b = s->tile_geom.b;
        s->tile_geom.tdim = make_uint3((s->dim.x + b.x - 1) / b.x, (s->dim.y + b.y - 1) / b.y, (s->dim.z + b.z - 1) / b.z);
        result = true;
        }
    if (c[0] > s->tile_Nmax)
        {
        s->tile_Nmax = round_up_u(c[0] + c[0] / 8 + 1, 8);
        tile_alloc_list(s);
        result = true;
        }
    if (result)
        {
        CUDA_OK(cudaMemsetAsync(s->tile_cond.p, 0, sizeof(unsigned int) * 2, s->stream));
        invalidate_graph(s);
        }
    return result;
    }

// host-driven rebuild (NeighborList::compute's overflow loop, hoomd/md/NeighborList.cc:226-244)
static void tile_nlist_build(b200_system* s)
    {
    const float rmax = s->rcut_max_max + s->r_buff;
    for (int d = 0; d < 3; d++)
        if (s->box.periodic[d] && s->box.L[d] <= rmax * 2.0f)
            throw std::runtime_error("nlist: Simulation box is too small! Particles would be interacting with themselves.");
    bool overflowed;
    do
        {
        tile_enqueue_rebuild(s, nullptr, s->pos.p);
        s->kernels_last_run += 7;
        cell_list_check_conditions(s);
        overflowed = tile_check_conditions(s);
        if (overflowed) s->num_overflows += 1;
        } while (overflowed);
    s->cl_first = false;
    s->cl_force = false;
    }
} // namespace b200
