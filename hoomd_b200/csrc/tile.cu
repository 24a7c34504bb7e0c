// CELL-TILE PATH: the MD step with neighbor positions served from shared memory instead of L1/L2 gathers.
//
// The CSR walk of fused_step.cu issues one fully divergent 16 B global gather per pair; on B200 that is bound by the L1
// tag stage (one wavefront per pair, ~70 % LSU utilisation at 1M particles) long before HBM or FP32.  This path removes
// the gathers:
//   * at every neighbor-list rebuild the particles are physically re-ordered by cell (row-major cell id, index order
//     inside a cell) -- the reference's CellList order (hoomd/CellList.cc:371-470) made the storage order, which also
//     subsumes the SFCPackUpdater's job (hoomd/SFCPackUpdater.cc:70-108: locality only, no observable effect);
//   * a CTA owns a block of b.x * b.y * b.z cells and stages those cells plus one halo layer of positions into shared
//     memory with bulk copies (cp.async.bulk + mbarrier: cells consecutive in x are consecutive in memory, so a halo row
//     is at most three contiguous runs) -- no per-element staging instructions;
//   * the neighbor list of an owned particle is stored as 16-bit SLOTS INTO THAT TILE, in chunks of 8 (chunk c of
//     particle g is the 16 B vector ell4[c * pitch + g]: 2 B per pair instead of 4 B, one fully coalesced 512 B load per
//     warp and 8 pairs, the next chunk in flight while this one is evaluated);
//   * 1, 2 or 4 lanes per owned particle read neighbor positions with LDS.128 and carry the particle from the pair sum
//     straight through step two / step one of the next step / the distance check in registers (md_step_body.cuh);
//     after the staging wait the kernel has no CTA barrier.
// The list holds exactly the pairs of NeighborListBinned (r <= r_cut + r_buff at build time, same cell grid), rebuilt
// on the same distance-check decisions, so forces differ from the CSR path only by the order of the FP32 sum.
#include "md_step_body.cuh"
#include "tile.h"
#include "bulk_copy.cuh"

namespace b200
{
constexpr int TILE_MAX_HALO = 128; // halo cells per tile: (b.x+2)(b.y+2)(b.z+2)
constexpr int TILE_MAX_RUNS = 16;  // owned (y,z) rows per tile: b.y * b.z

struct TileCtx
    {
    unsigned int hbeg[TILE_MAX_HALO + 1]; // first particle of halo cell h in the cell-sorted arrays
    unsigned int toff[TILE_MAX_HALO + 1]; // first tile slot of halo cell h (exclusive prefix of the cell sizes)
    unsigned int own_t0[TILE_MAX_RUNS];   // owned row r: first tile slot,
    unsigned int own_g0[TILE_MAX_RUNS];   //              first particle index,
    unsigned int own_off[TILE_MAX_RUNS + 1]; //           exclusive prefix of the owned counts
    unsigned int hxn, hyn, hzn, ownx, owny, ownz, n_own, total, wrap;
    };

// Tile coordinates = blockIdx (3-D grid).  Halo cell h = (hz*hyn + hy)*hxn + hx covers cell (x0-1+hx, y0-1+hy, z0-1+hz),
// wrapped on periodic axes, empty outside the grid otherwise.  Computed by warp 0; ends with a CTA barrier.
__device__ __forceinline__ void tile_setup(TileCtx& c, const TileGeom& g, const unsigned int* __restrict__ cell_start,
                                           uchar3 periodic)
    {
    if (threadIdx.x < 32)
        {
        const unsigned int x0 = blockIdx.x * g.b.x, y0 = blockIdx.y * g.b.y, z0 = blockIdx.z * g.b.z;
        const unsigned int ownx = min(g.b.x, g.dim.x - x0), owny = min(g.b.y, g.dim.y - y0),
                           ownz = min(g.b.z, g.dim.z - z0);
        const unsigned int hxn = ownx + 2, hyn = owny + 2, hzn = ownz + 2, nh = hxn * hyn * hzn;
        const unsigned int lane = threadIdx.x;
        unsigned int len[4], beg[4], sum = 0;
#pragma unroll
        for (int k = 0; k < 4; k++)
            {
            const unsigned int h = lane * 4 + k;
            len[k] = 0;
            beg[k] = 0;
            if (h < nh)
                {
                const unsigned int hx = h % hxn, hyz = h / hxn;
                int cx = (int)x0 - 1 + (int)hx, cy = (int)y0 - 1 + (int)(hyz % hyn), cz = (int)z0 - 1 + (int)(hyz / hyn);
                bool valid = true;
                if (cx < 0) { cx += (int)g.dim.x; valid &= periodic.x != 0; }
                else if (cx >= (int)g.dim.x) { cx -= (int)g.dim.x; valid &= periodic.x != 0; }
                if (cy < 0) { cy += (int)g.dim.y; valid &= periodic.y != 0; }
                else if (cy >= (int)g.dim.y) { cy -= (int)g.dim.y; valid &= periodic.y != 0; }
                if (cz < 0) { cz += (int)g.dim.z; valid &= periodic.z != 0; }
                else if (cz >= (int)g.dim.z) { cz -= (int)g.dim.z; valid &= periodic.z != 0; }
                if (valid)
                    {
                    const unsigned int cell = ((unsigned int)cz * g.dim.y + (unsigned int)cy) * g.dim.x + (unsigned int)cx;
                    beg[k] = cell_start[cell];
                    len[k] = cell_start[cell + 1] - beg[k];
                    }
                }
            sum += len[k];
            }
        unsigned int incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
            {
            const unsigned int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (unsigned int)o) incl += v;
            }
        unsigned int off = incl - sum;
#pragma unroll
        for (int k = 0; k < 4; k++)
            {
            c.hbeg[lane * 4 + k] = beg[k];
            c.toff[lane * 4 + k] = off;
            off += len[k];
            }
        if (lane == 31)
            {
            c.toff[TILE_MAX_HALO] = incl;
            c.total = incl;
            c.hxn = hxn; c.hyn = hyn; c.hzn = hzn;
            c.ownx = ownx; c.owny = owny; c.ownz = ownz;
            // the staged cells reach a periodic face (or the axis is too short for "two cells < L/2"): pairs of this
            // tile need minImage; elsewhere it is the identity and is skipped (CTA-uniform)
            const bool wx = periodic.x && (g.dim.x < 5 || x0 <= 1 || x0 + ownx + 1 >= g.dim.x);
            const bool wy = periodic.y && (g.dim.y < 5 || y0 <= 1 || y0 + owny + 1 >= g.dim.y);
            const bool wz = periodic.z && (g.dim.z < 5 || z0 <= 1 || z0 + ownz + 1 >= g.dim.z);
            c.wrap = (wx || wy || wz) ? 1u : 0u;
            }
        __syncwarp();
        // owned rows: lane ro = oz*owny + oy
        const unsigned int nro = owny * ownz;
        unsigned int cnt = 0, hrow = 0;
        if (lane < nro && nro <= TILE_MAX_RUNS)
            {
            hrow = ((lane / owny + 1) * hyn + (lane % owny + 1)) * hxn + 1;
            cnt = c.toff[hrow + ownx] - c.toff[hrow];
            }
        unsigned int oincl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
            {
            const unsigned int v = __shfl_up_sync(0xffffffffu, oincl, o);
            if (lane >= (unsigned int)o) oincl += v;
            }
        if (lane < nro && nro <= TILE_MAX_RUNS)
            {
            c.own_t0[lane] = c.toff[hrow];
            c.own_g0[lane] = c.hbeg[hrow];
            c.own_off[lane] = oincl - cnt;
            }
        if (lane == 31)
            {
            c.own_off[min(nro, (unsigned int)TILE_MAX_RUNS)] = oincl;
            c.n_own = oincl;
            }
        }
    __syncthreads();
    }

// Stage the halo block into shared memory with bulk copies (one per contiguous run: per (y,z) row the low halo cell,
// the owned cells and the high halo cell -- the end cells may be periodic images stored elsewhere).  Issued by the lanes
// of ONE warp; every thread of the CTA then waits on the mbarrier.
__device__ __forceinline__ void tile_stage_bulk(const TileCtx& c, float4* s_pos, const float4* __restrict__ pos,
                                                unsigned long long* bar)
    {
    const unsigned int lane = threadIdx.x & 31;
    if (lane == 0) mbar_arrive_expect_tx(bar, c.total * (unsigned int)sizeof(float4));
    __syncwarp();
    const unsigned int nseg = c.hyn * c.hzn * 3;
    for (unsigned int sg = lane; sg < nseg; sg += 32)
        {
        const unsigned int r = sg / 3, k = sg - 3 * r;
        const unsigned int h0 = r * c.hxn;
        const unsigned int hlo = h0 + (k == 0 ? 0u : k == 1 ? 1u : c.hxn - 1u);
        const unsigned int hhi = h0 + (k == 0 ? 1u : k == 1 ? c.hxn - 1u : c.hxn);
        const unsigned int t0 = c.toff[hlo], t1 = c.toff[hhi];
        if (t1 > t0) bulk_g2s(s_pos + t0, pos + c.hbeg[hlo], (t1 - t0) * (unsigned int)sizeof(float4), bar);
        }
    }

// owned particle p (0 .. n_own) -> its tile slot and its index in the particle arrays
__device__ __forceinline__ void tile_own(const TileCtx& c, unsigned int p, unsigned int& t, unsigned int& g,
                                         unsigned int& ro)
    {
    const unsigned int nro = c.owny * c.ownz;
    ro = 0;
    while (ro + 1 < nro && c.own_off[ro + 1] <= p) ro++;
    const unsigned int o = p - c.own_off[ro];
    t = c.own_t0[ro] + o;
    g = c.own_g0[ro] + o;
    }

__device__ __forceinline__ float4 tile_sentinel() { return make_float4(1.0e18f, 1.0e18f, 1.0e18f, 0.0f); }

// ---- re-order the particle data by cell ---------------------------------------------------------------------------------
// perm[e] = previous index of the particle now stored at e (the CellList's cell_idx); sorted positions = cell_xyzf
__global__ void __launch_bounds__(256) k_tile_gather(TileSortArgs a)
    {
    if (a.enable && !*a.enable) return;
    const unsigned int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= a.N) return;
    const unsigned int n = a.perm[e];
    a.vel_alt[e] = a.vel[n];
    const float3s* accel = (const float3s*)a.accel;
    ((float3s*)a.accel_alt)[e] = accel[n];
    const int3* image = (const int3*)a.image;
    ((int3*)a.image_alt)[e] = image[n];
    a.tag_alt[e] = a.tag[n];
    }

__global__ void __launch_bounds__(256) k_tile_copyback(TileSortArgs a)
    {
    if (a.enable && !*a.enable) return;
    const unsigned int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= a.N) return;
    const float4 p = a.sorted_pos[e];
    a.pos[e] = p;
    a.last_pos[e] = make_float4(p.x, p.y, p.z, 0.0f); // NeighborList::setLastUpdatedPos
    a.vel[e] = a.vel_alt[e];
    ((float3s*)a.accel)[e] = ((const float3s*)a.accel_alt)[e];
    ((int3*)a.image)[e] = ((const int3*)a.image_alt)[e];
    const unsigned int t = a.tag_alt[e];
    a.tag[e] = t;
    if (a.rtag) a.rtag[t] = e;
    }

cudaError_t launch_tile_sort(const TileSortArgs& a, cudaStream_t stream)
    {
    if (a.N == 0) return cudaSuccess;
    k_tile_gather<<<(a.N + 255) / 256, 256, 0, stream>>>(a);
    k_tile_copyback<<<(a.N + 255) / 256, 256, 0, stream>>>(a);
    return cudaGetLastError();
    }

static dim3 tile_grid(const TileGeom& g) { return dim3(g.tdim.x, g.tdim.y, g.tdim.z); }

// ---- build the tile-slot neighbor lists ------------------------------------------------------------------------------------
// One LANE per owned particle: it walks the candidates of the 9 (y,z) rows x 3 cells around its cell (3 consecutive cells
// of a row are one contiguous slot range) out of shared memory and appends the hits to its own row -- no ballots, no
// cross-lane compaction, rows in (row, slot) order.  The filter is NeighborListBinned's: r_cut(ti,tj) > 0 and
// dr^2 <= (r_cut + r_buff)^2 (hoomd/md/NeighborListBinned.cc:174-233, full storage, no body/diameter filters).
// A row is stored in chunks of 8 slots (one 16 B vector: chunk c of particle g at ell4[c * pitch + g]); the last chunk is
// padded with the sentinel slot (geom.cap: a position at 1e18).
__device__ __forceinline__ void tile_chunk_put(uint4& acc, unsigned int e, unsigned int t)
    {
    // entry e (0..7) of the chunk being filled; 64-bit halves so the shift amount stays a plain variable shift
    unsigned long long lo = ((unsigned long long)acc.y << 32) | acc.x, hi = ((unsigned long long)acc.w << 32) | acc.z;
    const unsigned long long v = (unsigned long long)t << ((e & 3u) * 16u);
    if (e < 4u) lo |= v; else hi |= v;
    acc = make_uint4((unsigned int)lo, (unsigned int)(lo >> 32), (unsigned int)hi, (unsigned int)(hi >> 32));
    }

template<bool MULTI_TYPE>
__global__ void __launch_bounds__(TILE_BLOCK) k_tile_build(const __grid_constant__ TileBuildArgs a)
    {
    if (a.enable && !*a.enable) return;
    extern __shared__ float4 s_pos[];
    __shared__ TileCtx c;
    __shared__ __align__(8) unsigned long long s_bar;
    if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    tile_setup(c, a.geom, a.cell_start, a.box.periodic);
    if (c.total > a.geom.cap || c.owny * c.ownz > TILE_MAX_RUNS)
        {
        if (threadIdx.x == 0) atomicMax(&a.conditions[1], c.total);
        return;
        }
    if (threadIdx.x < 32) tile_stage_bulk(c, s_pos, a.pos, &s_bar);
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const float rcut_00 = MULTI_TYPE ? 0.0f : a.r_cut[0];
    const float rlsq_00 = MULTI_TYPE ? 0.0f : a.r_listsq[0];
    const bool wrap = c.wrap != 0;
    const Box box = a.box;
    const unsigned int n_own = c.n_own;
    const unsigned int pad = a.geom.cap * 0x00010001u; // two sentinel slots
    mbar_wait(&s_bar, 0);
    for (unsigned int p = warp * 32 + lane; p < n_own; p += nwarps * 32)
        {
        unsigned int tp, gp, ro;
        tile_own(c, p, tp, gp, ro);
        const unsigned int oy = ro % c.owny, oz = ro / c.owny;
        // cell of p inside its row: hx in [1, ownx]
        const unsigned int hrow = ((oz + 1) * c.hyn + (oy + 1)) * c.hxn;
        unsigned int hx = 1;
        while (hx < c.ownx && c.toff[hrow + hx + 1] <= tp) hx++;
        const float4 pi4 = s_pos[tp];
        const unsigned int type_i = (unsigned int)__float_as_int(pi4.w);
        uint4* row = a.ell4 + gp;
        uint4 acc = make_uint4(0u, 0u, 0u, 0u);
        unsigned int n = 0;
        for (unsigned int dz = 0; dz < 3; dz++)
            for (unsigned int dy = 0; dy < 3; dy++)
                {
                const unsigned int h = ((oz + dz) * c.hyn + (oy + dy)) * c.hxn + hx - 1;
                const unsigned int lo = c.toff[h], hi = c.toff[h + 3];
                for (unsigned int t = lo; t < hi; t++)
                    {
                    const float4 cj = s_pos[t];
                    float rlsq = rlsq_00;
                    bool excluded = (t == tp);
                    if (MULTI_TYPE)
                        {
                        const unsigned int tj = (unsigned int)__float_as_int(cj.w);
                        excluded |= __ldg(&a.r_cut[type_i * a.ntypes + tj]) <= 0.0f;
                        rlsq = __ldg(&a.r_listsq[type_i * a.ntypes + tj]);
                        }
                    else
                        excluded |= rcut_00 <= 0.0f;
                    float3 dx = make_float3(__fsub_rn(pi4.x, cj.x), __fsub_rn(pi4.y, cj.y), __fsub_rn(pi4.z, cj.z));
                    if (wrap) dx = min_image_exact(dx, box);
                    if (!excluded && dot_exact(dx) <= rlsq)
                        {
                        if (n < a.Nmax)
                            {
                            tile_chunk_put(acc, n & 7u, t);
                            if ((n & 7u) == 7u)
                                {
                                row[(size_t)(n >> 3) * a.pitch] = acc;
                                acc = make_uint4(0u, 0u, 0u, 0u);
                                }
                            }
                        n++;
                        }
                    }
                }
        const unsigned int kept = min(n, a.Nmax);
        if (kept & 7u)
            {
            for (unsigned int e = kept & 7u; e < 8u; e++) tile_chunk_put(acc, e, a.geom.cap);
            row[(size_t)(kept >> 3) * a.pitch] = acc;
            }
        (void)pad;
        a.nn[gp] = kept;
        if (n > a.Nmax) atomicMax(&a.conditions[0], n);
        }
    }

cudaError_t launch_tile_build(const TileBuildArgs& a, cudaStream_t stream)
    {
    const size_t smem = sizeof(float4) * (a.geom.cap + 1);
    if (a.ntypes > 1)
        {
        cudaFuncSetAttribute(k_tile_build<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_tile_build<true><<<tile_grid(a.geom), TILE_BLOCK, smem, stream>>>(a);
        }
    else
        {
        cudaFuncSetAttribute(k_tile_build<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_tile_build<false><<<tile_grid(a.geom), TILE_BLOCK, smem, stream>>>(a);
        }
    return cudaGetLastError();
    }

__global__ void k_tile_post(StepState* st, const unsigned int* __restrict__ conditions, unsigned int Nmax)
    {
    if (!st->need_rebuild) return;
    if (conditions[0] > Nmax || conditions[1]) st->overflow = 1u;
    }

cudaError_t launch_tile_post(StepState* st, const unsigned int* conditions, unsigned int Nmax, cudaStream_t stream)
    {
    k_tile_post<<<1, 1, 0, stream>>>(st, conditions, Nmax);
    return cudaGetLastError();
    }

// ---- tile slots -> particle indices (the list in NeighborList's own terms, for read-back) ------------------------------------
__global__ void __launch_bounds__(TILE_BLOCK) k_tile_to_csr(const __grid_constant__ TileBuildArgs a,
                                                            unsigned int* __restrict__ out_n,
                                                            unsigned int* __restrict__ out_head,
                                                            unsigned int* __restrict__ out_list)
    {
    __shared__ TileCtx c;
    tile_setup(c, a.geom, a.cell_start, a.box.periodic);
    const unsigned int nh = c.hxn * c.hyn * c.hzn;
    for (unsigned int p = threadIdx.x; p < c.n_own; p += blockDim.x)
        {
        unsigned int tp, gp, ro;
        tile_own(c, p, tp, gp, ro);
        const unsigned int n = a.nn[gp];
        out_n[gp] = n;
        out_head[gp] = gp * a.Nmax;
        for (unsigned int k = 0; k < n; k++)
            {
            const unsigned int t = ((const unsigned short*)(a.ell4 + (size_t)(k >> 3) * a.pitch + gp))[k & 7u];
            unsigned int lo = 0, hi = nh; // last h with toff[h] <= t
            while (hi - lo > 1)
                {
                const unsigned int mid = (lo + hi) >> 1;
                if (c.toff[mid] <= t) lo = mid; else hi = mid;
                }
            out_list[(size_t)gp * a.Nmax + k] = c.hbeg[lo] + (t - c.toff[lo]);
            }
        }
    }

cudaError_t launch_tile_to_csr(const TileBuildArgs& a, unsigned int* out_n, unsigned int* out_head,
                               unsigned int* out_list, cudaStream_t stream)
    {
    k_tile_to_csr<<<tile_grid(a.geom), TILE_BLOCK, 0, stream>>>(a, out_n, out_head, out_list);
    return cudaGetLastError();
    }

// ---- the MD step over cell tiles ---------------------------------------------------------------------------------------------
// TPP lanes per owned particle (1 for large systems: nothing is left of the divergent-gather argument for more; 2 or 4
// when there are too few particles to fill the SMs): lane `sub` takes the row's chunks sub, sub + TPP, ... -- one 16 B
// vector load per 8 pairs, the next chunk in flight while this one is evaluated out of shared memory.  The particle then
// goes straight through step two / step one / the distance check in registers: after the staging wait there is no CTA
// barrier until the ticket.
template<class Evaluator, int SHIFT_MODE, bool VIRIAL, bool MULTI_TYPE, bool WRAP>
__device__ __forceinline__ void tile_eval4(const float4 pi4, unsigned int ta, unsigned int tb, unsigned int typei,
                                           const float4* __restrict__ s_pos, const Box& box,
                                           const PairTables<Evaluator>& T, unsigned int ntypes, float4& F, float (&vir)[6])
    {
    float4 pj[4];
    pj[0] = s_pos[ta & 0xffffu];
    pj[1] = s_pos[ta >> 16];
    pj[2] = s_pos[tb & 0xffffu];
    pj[3] = s_pos[tb >> 16];
#pragma unroll
    for (int u = 0; u < 4; u++)
        pair_one<Evaluator, SHIFT_MODE, VIRIAL, MULTI_TYPE, WRAP>(pi4, pj[u], typei, box, T, ntypes, F.x, F.y, F.z, F.w,
                                                                  vir[0], vir[1], vir[2], vir[3], vir[4], vir[5]);
    }

template<class Evaluator, int TPP, int SHIFT_MODE, bool VIRIAL, bool MULTI_TYPE, bool WRAP>
__device__ __forceinline__ void tile_walk(const float4 pi4, uint4 cur, unsigned int nc, unsigned int sub,
                                          const uint4* __restrict__ row, unsigned int pitch,
                                          const float4* __restrict__ s_pos, const Box& box,
                                          const PairTables<Evaluator>& T, unsigned int ntypes, float4& F, float (&vir)[6])
    {
    const unsigned int typei = (unsigned int)__float_as_int(pi4.w);
    for (unsigned int ch = sub; ch < nc; ch += TPP)
        {
        uint4 nxt = cur;
        if (ch + TPP < nc) nxt = __ldcs(row + (size_t)(ch + TPP) * pitch);
        tile_eval4<Evaluator, SHIFT_MODE, VIRIAL, MULTI_TYPE, WRAP>(pi4, cur.x, cur.y, typei, s_pos, box, T, ntypes, F, vir);
        tile_eval4<Evaluator, SHIFT_MODE, VIRIAL, MULTI_TYPE, WRAP>(pi4, cur.z, cur.w, typei, s_pos, box, T, ntypes, F, vir);
        cur = nxt;
        }
    }

template<class Evaluator, int TPP, int SHIFT_MODE, bool VIRIAL, bool MULTI_TYPE>
__global__ void __launch_bounds__(TILE_BLOCK* TPP) k_md_step_tile(const __grid_constant__ MdStepArgs<Evaluator> a)
    {
    constexpr unsigned int PPB = TILE_BLOCK; // owned particles per round of the CTA
    extern __shared__ float4 s_dyn[];        // [cap + 1] staged positions (+ sentinel), then the per-type-pair tables
    __shared__ TileCtx c;
    __shared__ __align__(8) unsigned long long s_bar;
    float4* s_pos = s_dyn;
    StepState* st = a.st;
    unsigned int steps_left = 1;
    bool last = true, store_force = true;
    unsigned int ts2 = 0;
    if (!a.force_only)
        {
        steps_left = st->steps_left;
        if (steps_left == 0 || st->abort) return; // padding launch of an odd-length chunk (grid-uniform)
        last = steps_left == 1;
        ts2 = st->ts_step_two;
        store_force = last || st->store_force;
        }
    if (threadIdx.x == 0)
        {
        mbar_init(&s_bar, 1);
        s_pos[a.tile.cap] = tile_sentinel();
        }
    PairTables<Evaluator> T;
    pair_tables_load<Evaluator, SHIFT_MODE, MULTI_TYPE>(T, reinterpret_cast<unsigned char*>(s_dyn + a.tile.cap + 1),
                                                        a.params, a.rcutsq, a.ronsq, a.ntypes);
    tile_setup(c, a.tile, a.cell_start, a.box.periodic);
    const bool ok = c.total <= a.tile.cap && c.owny * c.ownz <= TILE_MAX_RUNS; // else: the build flagged the overflow
    bool moved = false;
    if (ok)
        {
        if (threadIdx.x < 32) tile_stage_bulk(c, s_pos, a.pos_in, &s_bar);
        const unsigned int slot = threadIdx.x / TPP, sub = threadIdx.x % TPP;
        const unsigned int n_own = c.n_own;
        const bool wrap = c.wrap != 0;
        // rounds are warp-uniform: a warp holds 32 / TPP consecutive owned particles
        for (unsigned int p0 = 0; p0 < n_own; p0 += PPB)
            {
            const unsigned int p = p0 + slot;
            const bool active = p < n_own;
            unsigned int tp = 0, gp = 0, ro;
            if (active) tile_own(c, p, tp, gp, ro);
            // the row's first chunk and its length do not depend on the staged positions: get them in flight first
            const uint4* row = a.ell4 + gp;
            uint4 cur = make_uint4(0u, 0u, 0u, 0u);
            unsigned int n = 0;
            if (active)
                {
                if (sub * 8u < a.ell_nmax) cur = __ldcs(row + (size_t)sub * a.ell_pitch);
                n = a.n_neigh[gp];
                }
            if (p0 == 0) mbar_wait(&s_bar, 0);
            const float4 pi4 = active ? s_pos[tp] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            const unsigned int nc = (n + 7u) >> 3;
            float4 F = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            float vir[6] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
            if (wrap)
                tile_walk<Evaluator, TPP, SHIFT_MODE, VIRIAL, MULTI_TYPE, true>(pi4, cur, nc, sub, row, a.ell_pitch, s_pos,
                                                                                a.box, T, a.ntypes, F, vir);
            else
                tile_walk<Evaluator, TPP, SHIFT_MODE, VIRIAL, MULTI_TYPE, false>(pi4, cur, nc, sub, row, a.ell_pitch, s_pos,
                                                                                 a.box, T, a.ntypes, F, vir);
            if (TPP > 1)
                {
                F.x = group_reduce<TPP>(F.x);
                F.y = group_reduce<TPP>(F.y);
                F.z = group_reduce<TPP>(F.z);
                F.w = group_reduce<TPP>(F.w);
                if (VIRIAL)
                    {
#pragma unroll
                    for (int k = 0; k < 6; k++) vir[k] = group_reduce<TPP>(vir[k]);
                    }
                }
            if (active && sub == 0)
                {
                if (a.force_only)
                    {
                    a.net_force[gp] = F;
                    if (VIRIAL)
                        {
#pragma unroll
                        for (int k = 0; k < 6; k++) a.virial[k * a.virial_pitch + gp] = vir[k];
                        }
                    }
                else
                    {
                    double mv2_unused = 0.0; // the NVT thermostat runs on the CSR path only (engine: add_nvt)
                    moved |= md_step_integrate<Evaluator, VIRIAL>(a, gp, F, pi4, vir, ts2, last, store_force, mv2_unused);
                    }
                }
            }
        if (n_own == 0) mbar_wait(&s_bar, 0); // do not exit with the bulk copies in flight
        }
    if (!a.force_only) md_step_finish(a, st, moved, steps_left, last);
    }

template<class Evaluator, int TPP, int SHIFT_MODE, bool VIRIAL>
static cudaError_t md_tile_mt(const MdStepArgs<Evaluator>& a, cudaStream_t stream)
    {
    size_t smem = sizeof(float4) * (a.tile.cap + 1);
    if (a.ntypes > 1)
        {
        smem += (size_t)a.ntypes * a.ntypes * (sizeof(typename Evaluator::param_type) + 2 * sizeof(float));
        cudaFuncSetAttribute(k_md_step_tile<Evaluator, TPP, SHIFT_MODE, VIRIAL, true>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_md_step_tile<Evaluator, TPP, SHIFT_MODE, VIRIAL, true><<<tile_grid(a.tile), TILE_BLOCK * TPP, smem, stream>>>(a);
        }
    else
        {
        cudaFuncSetAttribute(k_md_step_tile<Evaluator, TPP, SHIFT_MODE, VIRIAL, false>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_md_step_tile<Evaluator, TPP, SHIFT_MODE, VIRIAL, false><<<tile_grid(a.tile), TILE_BLOCK * TPP, smem, stream>>>(a);
        }
    return cudaGetLastError();
    }

template<class Evaluator, int TPP>
static cudaError_t md_tile_tpp(const MdStepArgs<Evaluator>& a, unsigned int shift_mode, bool v, cudaStream_t stream)
    {
    switch (shift_mode)
        {
        case B200_SHIFT_NONE:
            return v ? md_tile_mt<Evaluator, TPP, B200_SHIFT_NONE, true>(a, stream)
                     : md_tile_mt<Evaluator, TPP, B200_SHIFT_NONE, false>(a, stream);
        case B200_SHIFT_SHIFT:
            return v ? md_tile_mt<Evaluator, TPP, B200_SHIFT_SHIFT, true>(a, stream)
                     : md_tile_mt<Evaluator, TPP, B200_SHIFT_SHIFT, false>(a, stream);
        case B200_SHIFT_XPLOR:
            return v ? md_tile_mt<Evaluator, TPP, B200_SHIFT_XPLOR, true>(a, stream)
                     : md_tile_mt<Evaluator, TPP, B200_SHIFT_XPLOR, false>(a, stream);
        default:
            return cudaErrorInvalidValue;
        }
    }

cudaError_t launch_md_step_tile_lj(const b200_pair_args& pa, const float* d_params, const FusedStepArgs& fa,
                                   const TileStepArgs& ta, const float4* pos_in, float4* pos_out, int copy_pos_on_last,
                                   StepState* st, unsigned long long cond_handle, int use_cond, cudaStream_t stream)
    {
    if (fa.N == 0) return cudaSuccess;
    if (!box_is_orthorhombic(pa.box)) return cudaErrorNotSupported;
    MdStepArgs<EvaluatorPairLJ> a;
    fill_md_step_args(a, pa, d_params, fa, pos_in, pos_out, copy_pos_on_last, st, cond_handle, use_cond);
    a.n_neigh = ta.nn;
    a.tile = ta.geom;
    a.cell_start = ta.cell_start;
    a.ell4 = ta.ell4;
    a.ell_pitch = ta.pitch;
    a.ell_nmax = ta.Nmax;
    a.force_only = ta.force_only;
    const bool v = pa.compute_virial != 0;
    switch (ta.tpp)
        {
        case 1: return md_tile_tpp<EvaluatorPairLJ, 1>(a, pa.shift_mode, v, stream);
        case 2: return md_tile_tpp<EvaluatorPairLJ, 2>(a, pa.shift_mode, v, stream);
        case 4: return md_tile_tpp<EvaluatorPairLJ, 4>(a, pa.shift_mode, v, stream);
        default: return cudaErrorInvalidValue;
        }
    }
} // namespace b200
