// NeighborListBinned for sm_100a.  One warp per particle walks the <=27 adjacent cells as <=9+ contiguous ranges of the
// cell-sorted array (x-adjacent cells are adjacent in memory), 32 candidates per coalesced 512 B load, ballot/popc
// compaction into the CSR list.  Membership arithmetic and entry order are those of the CPU oracle
// hoomd/md/NeighborListBinned.cc:144-236 (see include/hoomd_b200.h).
#include <cmath>
#include <algorithm>
#include "common.cuh"
#include "kernels.h"
#include "bulk_copy.cuh"
#include <cstdlib>

namespace b200
{
// sorted, de-duplicated {c-1, c, c+1} wrapped into [0,m) (periodic) or clipped (non-periodic: the reference wraps these
// too, hoomd/CellList.cc:399-411, but a wrapped cell of a non-periodic direction is >= one cell width away and can never
// pass the distance test, so skipping it changes nothing)
__device__ __forceinline__ int axis_set(int c, int m, bool periodic, int out[3])
    {
    int n = 0;
#pragma unroll
    for (int d = -1; d <= 1; d++)
        {
        int w = c + d;
        if (w < 0) { if (!periodic) continue; w += m; if (w < 0) w += m; }
        if (w >= m) { if (!periodic) continue; w -= m; if (w >= m) w -= m; }
        // insert into the sorted unique list
        bool dup = false;
        for (int k = 0; k < n; k++) dup |= (out[k] == w);
        if (dup) continue;
        int k = n;
        while (k > 0 && out[k - 1] > w) { out[k] = out[k - 1]; k--; }
        out[k] = w;
        n++;
        }
    return n;
    }

// the same set in registers (no local array, no insertion loop): c and its wrapped / clipped neighbors, duplicates dropped
// (dim 1 or 2), ascending, absent entries at the end
__device__ __forceinline__ int axis_set3(int c, int m, bool periodic, int& o0, int& o1, int& o2)
    {
    constexpr int ABSENT = 0x3fffffff;
    int lo = c - 1, hi = c + 1;
    bool vlo = true, vhi = true;
    if (lo < 0) { vlo = periodic; lo += m; }
    if (hi >= m) { vhi = periodic; hi -= m; }
    if (lo == c) vlo = false;                // m == 1
    if (hi == c || (vlo && hi == lo)) vhi = false; // m == 1, m == 2
    const int a = vlo ? lo : ABSENT, b = vhi ? hi : ABSENT;
    const int mn = min(min(a, b), c), mx = max(max(a, b), c);
    o0 = mn;
    o2 = mx;
    o1 = a + b + c - mn - mx; // < 2^31: at most two ABSENT terms
    return 1 + (vlo ? 1 : 0) + (vhi ? 1 : 0);
    }

// on overflow the true count goes to `conditions` (the host grows Nmax and rebuilds, NeighborList.cc:226-244)
__device__ __forceinline__ void nlist_report_overflow(unsigned int* conditions, unsigned int type_i, unsigned int n,
                                                      const OverflowHook& hook)
    {
    atomicMax(&conditions[type_i], n);
    if (hook.overflow) *hook.overflow = 1u;
    if (hook.steps_left) *hook.steps_left = 0u;
    }

template<bool MULTI_TYPE>
__global__ void __launch_bounds__(128) k_nlist_binned(unsigned int* __restrict__ nlist,
                                                      unsigned int* __restrict__ n_neigh,
                                                      float4* __restrict__ last_pos,
                                                      unsigned int* __restrict__ conditions,
                                                      const unsigned int* __restrict__ Nmax,
                                                      const unsigned int* __restrict__ head_list,
                                                      const float4* __restrict__ pos, unsigned int N,
                                                      const unsigned int* __restrict__ cell_start,
                                                      const float4* __restrict__ cell_xyzf,
                                                      const unsigned int* __restrict__ cell_type, uint3 dim, Box box,
                                                      float3 ghost_width, const float* __restrict__ r_cut,
                                                      const float* __restrict__ r_listsq, unsigned int ntypes,
                                                      int half_storage, const unsigned int* __restrict__ enable,
                                                      OverflowHook hook)
    {
    pdl_wait();
    if (enable && !*enable) return;
    const unsigned int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned int lane = threadIdx.x & 31;
    if (i >= N) return;
    const float4 pi4 = pos[i];
    const float3 pi = make_float3(pi4.x, pi4.y, pi4.z);
    const unsigned int type_i = (unsigned int)__float_as_int(pi4.w);
    const unsigned int Nmax_i = Nmax[type_i];
    const unsigned int head_i = head_list[i];
    // single-type fast path keeps the two table entries in registers
    const float rcut_00 = MULTI_TYPE ? 0.0f : r_cut[0];
    const float rlsq_00 = MULTI_TYPE ? 0.0f : r_listsq[0];

    int3 b;
    unsigned int n = 0;
    if (particle_bin(pi, box, ghost_width, dim, b))
        {
        int xs[3], ys[3], zs[3];
        const int nx = axis_set(b.x, dim.x, box.periodic.x, xs);
        const int ny = axis_set(b.y, dim.y, box.periodic.y, ys);
        const int nz = axis_set(b.z, dim.z, box.periodic.z, zs);
        const unsigned int lanemask_lt = (1u << lane) - 1u;
        for (int c = 0; c < nz; c++)
            for (int q = 0; q < ny; q++)
                {
                const unsigned int row = ((unsigned int)zs[c] * dim.y + (unsigned int)ys[q]) * dim.x;
                int r0 = 0;
                while (r0 < nx)
                    {
                    int r1 = r0;
                    while (r1 + 1 < nx && xs[r1 + 1] == xs[r1] + 1) r1++;
                    const unsigned int beg = cell_start[row + xs[r0]];
                    const unsigned int end = cell_start[row + xs[r1] + 1];
                    for (unsigned int base = beg; base < end; base += 32)
                        {
                        const unsigned int e = base + lane;
                        bool hit = false;
                        unsigned int j = 0;
                        if (e < end)
                            {
                            const float4 cj = cell_xyzf[e];
                            j = (unsigned int)__float_as_int(cj.w);
                            float rc, rlsq;
                            if (MULTI_TYPE)
                                {
                                const unsigned int tj = cell_type[e];
                                rc = __ldg(&r_cut[type_i * ntypes + tj]);
                                rlsq = __ldg(&r_listsq[type_i * ntypes + tj]);
                                }
                            else
                                {
                                rc = rcut_00;
                                rlsq = rlsq_00;
                                }
                            const bool excluded = (j == i) || (rc <= 0.0f);
                            float3 dx = make_float3(__fsub_rn(pi.x, cj.x), __fsub_rn(pi.y, cj.y), __fsub_rn(pi.z, cj.z));
                            dx = min_image_exact(dx, box);
                            const float drsq = dot_exact(dx);
                            hit = !excluded && (drsq <= rlsq) && (!half_storage || i < j);
                            }
                        const unsigned int mask = __ballot_sync(0xffffffffu, hit);
                        if (hit)
                            {
                            const unsigned int k = n + __popc(mask & lanemask_lt);
                            if (k < Nmax_i) nlist[head_i + k] = j;
                            }
                        n += __popc(mask);
                        }
                    r0 = r1 + 1;
                    }
                }
        }
    if (lane == 0)
        {
        // on overflow the true count goes to `conditions` (the host grows Nmax and rebuilds, NeighborList.cc:226-244);
        // the stored count is clamped so that a consumer running before the retry never reads past this segment
        n_neigh[i] = min(n, Nmax_i);
        if (n > Nmax_i) nlist_report_overflow(conditions, type_i, n, hook);
        if (last_pos) last_pos[i] = make_float4(pi.x, pi.y, pi.z, 0.0f);
        }
    }

// ---- tiled build: one CTA per cell, the <=27 adjacent cells staged through shared memory -----------------------------
// Every particle of a cell tests the same candidates, so the CTA copies them ONCE (coalesced 16 B loads, in the CPU
// loop's order: neighbor cells by ascending id, occupants by ascending particle index) and its warps then take the
// cell's particles round-robin, each lane striding the staged tile: 32 candidates per ballot, popc compaction.
// Per particle this moves 27*occ*16/occ = 432 B through L2 instead of 27*occ*16 = 3.5 kB (rho = 0.2), and removes the
// per-particle dependent cell_start -> xyzf chain.  Membership arithmetic is unchanged (bit-exact vs the oracle).
#ifndef B200_NL_BLOCK
#define B200_NL_BLOCK 128
#endif
constexpr int NL_BLOCK = B200_NL_BLOCK; // threads per cell (A/B: -DB200_NL_BLOCK=256)
constexpr int NL_CAP = 1024; // staged candidates per pass (16 kB + 4 kB types); denser neighborhoods take several passes

// row[k] = v with the address formed from the row pointer held in registers (the compiler otherwise rebuilds
// nlist + head + k from the kernel parameter with five predicated 32-bit instructions per store: 16 % of the scan loop).
// No other access to the list in these kernels, so the store needs no ordering against the surrounding loads.
__device__ __forceinline__ void row_store(unsigned int* row, unsigned int k, unsigned int v)
    {
    unsigned long long addr;
    asm("mad.wide.u32 %0, %1, 4, %2;" : "=l"(addr) : "r"(k), "l"(__cvta_generic_to_global(row)));
    asm volatile("st.global.u32 [%0], %1;" ::"l"(addr), "r"(v));
    }

// One warp tests every staged candidate against particle i: 2 x 32 candidates per trip, ballot + popc compaction in the
// staged (= oracle) order.  The tile is padded to a multiple of 64 with far-away sentinels, so there is no bounds test.
template<bool MULTI_TYPE, bool WRAP, bool HALF, int U = 2>
__device__ __forceinline__ unsigned int nlist_scan_tile(unsigned int n, const float4 pi4, unsigned int i,
                                                        unsigned int type_i, unsigned int Nmax_i,
                                                        unsigned int* __restrict__ row, const float4* s_cand,
                                                        const unsigned int* s_type, unsigned int cnt_padded,
                                                        unsigned int lane, const Box& box, bool wx, bool wy, bool wz,
                                                        float rcut_00, float rlsq_00, const float* __restrict__ r_cut,
                                                        const float* __restrict__ r_listsq, unsigned int ntypes)
    {
    const unsigned int lanemask_lt = (1u << lane) - 1u;
    for (unsigned int base = 0; base < cnt_padded; base += 32 * U)
        {
        bool hit[U];
        unsigned int jj[U];
#pragma unroll
        for (int u = 0; u < U; u++)
            {
            const unsigned int t = base + 32 * u + lane;
            const float4 cj = s_cand[t];
            const unsigned int j = (unsigned int)__float_as_int(cj.w);
            float rlsq = rlsq_00;
            bool excluded = (j == i);
            if (MULTI_TYPE)
                {
                const unsigned int tj = s_type[t];
                excluded |= __ldg(&r_cut[type_i * ntypes + tj]) <= 0.0f;
                rlsq = __ldg(&r_listsq[type_i * ntypes + tj]);
                }
            else
                excluded |= rcut_00 <= 0.0f;
            float3 dx = make_float3(__fsub_rn(pi4.x, cj.x), __fsub_rn(pi4.y, cj.y), __fsub_rn(pi4.z, cj.z));
            if (WRAP) dx = min_image_exact_sel(dx, box, wx, wy, wz);
            const float drsq = dot_exact(dx);
            hit[u] = !excluded && (drsq <= rlsq) && (!HALF || i < j);
            jj[u] = j;
            }
#pragma unroll
        for (int u = 0; u < U; u++)
            {
            const unsigned int mask = __ballot_sync(0xffffffffu, hit[u]);
            const unsigned int k = n + __popc(mask & lanemask_lt);
            if (hit[u] && k < Nmax_i) row_store(row, k, jj[u]);
            n += __popc(mask);
            }
        }
    return n;
    }

template<bool MULTI_TYPE, bool HALF>
__global__ void __launch_bounds__(NL_BLOCK) k_nlist_tiled(unsigned int* __restrict__ nlist,
                                                          unsigned int* __restrict__ n_neigh,
                                                          float4* __restrict__ last_pos,
                                                          unsigned int* __restrict__ conditions,
                                                          const unsigned int* __restrict__ Nmax,
                                                          const unsigned int* __restrict__ head_list, unsigned int N,
                                                          const unsigned int* __restrict__ cell_start,
                                                          const float4* __restrict__ cell_xyzf,
                                                          const unsigned int* __restrict__ cell_type, uint3 dim,
                                                          Box box, const float* __restrict__ r_cut,
                                                          const float* __restrict__ r_listsq, unsigned int ntypes,
                                                          const unsigned int* __restrict__ enable, OverflowHook hook)
    {
    pdl_wait();
    if (enable && !*enable) return;
    extern __shared__ float4 s_cand[];
    unsigned int* s_type = reinterpret_cast<unsigned int*>(s_cand + NL_CAP);
    __shared__ unsigned int s_beg[28], s_off[28]; // per adjacent cell: first entry, exclusive prefix of the lengths
    __shared__ unsigned int s_nr;

    // grid = (dim.x, dim.y, dim.z): the cell coordinates come for free
    const unsigned int bx = blockIdx.x, by = blockIdx.y, bz = blockIdx.z;
    const unsigned int cell = (bz * dim.y + by) * dim.x + bx;
    const unsigned int own_beg = cell_start[cell], own_end = cell_start[cell + 1];
    if (own_beg == own_end) return; // uniform per CTA
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    if (warp == 0)
        {
        // (the register-only axis_set3 of the warp-per-cell kernel measured 1-2 us SLOWER here at 64k: 44 registers, or 40 with
        // spills, against 40 with this 40-byte stack frame)
        int xs[3], ys[3], zs[3];
        const int nx = axis_set(bx, dim.x, box.periodic.x, xs);
        const int ny = axis_set(by, dim.y, box.periodic.y, ys);
        const int nz = axis_set(bz, dim.z, box.periodic.z, zs);
        const int nr = nx * ny * nz;
        unsigned int beg = 0, len = 0;
        if ((int)lane < nr)
            {
            // (c, q, r) in lexicographic order == ascending cell id (id = (z*dimy + y)*dimx + x)
            const int r = lane % nx, q = (lane / nx) % ny, c = lane / (nx * ny);
            const int x = (r == 0) ? xs[0] : (r == 1) ? xs[1] : xs[2];
            const int y = (q == 0) ? ys[0] : (q == 1) ? ys[1] : ys[2];
            const int z = (c == 0) ? zs[0] : (c == 1) ? zs[1] : zs[2];
            const unsigned int nc = ((unsigned int)z * dim.y + (unsigned int)y) * dim.x + (unsigned int)x;
            beg = cell_start[nc];
            len = cell_start[nc + 1] - beg;
            }
        unsigned int incl = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
            {
            const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
            if ((int)lane >= o) incl += t;
            }
        if (lane < 28)
            {
            s_beg[lane] = beg;
            s_off[lane] = incl - len;
            }
        if ((int)lane == nr - 1)
            {
            s_off[nr] = incl;
            s_nr = nr;
            }
        }
    __syncthreads();
    const unsigned int nr = s_nr;
    const unsigned int total = s_off[nr];
    const float rcut_00 = MULTI_TYPE ? 0.0f : r_cut[0];
    const float rlsq_00 = MULTI_TYPE ? 0.0f : r_listsq[0];
    // A candidate of an interior cell is less than two cell widths away along every axis; with >= 5 cells per axis that is
    // inside [lo, hi) and minImage is the identity: only cells on a periodic face pay for it (CTA-uniform).
    const bool wx = cell_needs_image(box.iperiodic.x, box.periodic.x, box.iforce.x, box.ighost.x, dim.x, bx);
    const bool wy = cell_needs_image(box.iperiodic.y, box.periodic.y, box.iforce.y, box.ighost.y, dim.y, by);
    const bool wz = cell_needs_image(box.iperiodic.z, box.periodic.z, box.iforce.z, box.ighost.z, dim.z, bz);
    const bool wrap_any = wx || wy || wz;

    for (unsigned int pass = 0; pass < total; pass += NL_CAP)
        {
        const unsigned int cnt = min((unsigned int)NL_CAP, total - pass);
        const unsigned int cnt_padded = (cnt + 31u) & ~31u; // NL_CAP is a multiple of 64
        const unsigned int cnt64 = cnt_padded & ~63u;
        if (pass) __syncthreads(); // everyone is done with the previous tile
        for (unsigned int t = threadIdx.x; t < cnt_padded; t += NL_BLOCK)
            {
            if (t < cnt)
                {
                const unsigned int g = pass + t;
                // adjacent cell holding staged entry g: last r with s_off[r] <= g (<= 27 ranges: 5-step bisection; a half
                // warp per adjacent cell instead needs fewer instructions but measured 6 % SLOWER at 1M -- its loads
                // hang on three dependent shared-memory reads and leave half the lanes idle)
                unsigned int lo = 0, hi = nr;
                while (hi - lo > 1)
                    {
                    const unsigned int mid = (lo + hi) >> 1;
                    if (s_off[mid] <= g) lo = mid; else hi = mid;
                    }
                const unsigned int e = s_beg[lo] + (g - s_off[lo]);
                s_cand[t] = cell_xyzf[e];
                if (MULTI_TYPE) s_type[t] = cell_type[e];
                }
            else
                {
                // sentinel: dr^2 overflows to +inf, never <= r_list^2; index 0xffffffff is nobody
                s_cand[t] = make_float4(1.0e30f, 1.0e30f, 1.0e30f, __int_as_float(-1));
                if (MULTI_TYPE) s_type[t] = 0u;
                }
            }
        __syncthreads();

        for (unsigned int p = own_beg + warp; p < own_end; p += NL_BLOCK / 32)
            {
            const float4 pi4 = cell_xyzf[p];
            const unsigned int i = (unsigned int)__float_as_int(pi4.w);
            if (i >= N) continue; // ghosts are candidates only
            const unsigned int type_i = MULTI_TYPE ? cell_type[p] : 0u;
            const unsigned int Nmax_i = Nmax[type_i];
            unsigned int* row = nlist + head_list[i];
            unsigned int n = pass ? n_neigh[i] : 0u; // later passes continue the count this warp stored itself
            // 64 candidates per trip, then at most one 32-wide trip: the tile is padded to 32, not to 64 (a neighborhood
            // holds 216 +- 15 candidates: 6 % fewer distance tests)
            if (wrap_any)
                {
                n = nlist_scan_tile<MULTI_TYPE, true, HALF>(n, pi4, i, type_i, Nmax_i, row, s_cand, s_type, cnt64, lane, box,
                                                            wx, wy, wz, rcut_00, rlsq_00, r_cut, r_listsq, ntypes);
                if (cnt_padded != cnt64)
                    n = nlist_scan_tile<MULTI_TYPE, true, HALF, 1>(n, pi4, i, type_i, Nmax_i, row, s_cand + cnt64,
                                                                   s_type + cnt64, 32u, lane, box, wx, wy, wz, rcut_00,
                                                                   rlsq_00, r_cut, r_listsq, ntypes);
                }
            else
                {
                n = nlist_scan_tile<MULTI_TYPE, false, HALF>(n, pi4, i, type_i, Nmax_i, row, s_cand, s_type, cnt64, lane, box,
                                                             wx, wy, wz, rcut_00, rlsq_00, r_cut, r_listsq, ntypes);
                if (cnt_padded != cnt64)
                    n = nlist_scan_tile<MULTI_TYPE, false, HALF, 1>(n, pi4, i, type_i, Nmax_i, row, s_cand + cnt64,
                                                                    s_type + cnt64, 32u, lane, box, wx, wy, wz, rcut_00,
                                                                    rlsq_00, r_cut, r_listsq, ntypes);
                }
            if (lane == 0)
                {
                if (pass + NL_CAP < total)
                    n_neigh[i] = n; // running count (unclamped) for the next pass
                else
                    {
                    n_neigh[i] = min(n, Nmax_i);
                    if (n > Nmax_i) nlist_report_overflow(conditions, type_i, n, hook);
                    if (last_pos) last_pos[i] = make_float4(pi4.x, pi4.y, pi4.z, 0.0f);
                    }
                }
            __syncwarp();
            }
        }
    }

// ---- warp-per-cell build for sparse cells --------------------------------------------------------------------------------
// With ~3 particles per cell (bead-spring melt: rho 0.85, cell width 1.52) a 128-thread CTA per cell stages 81 candidates,
// keeps one warp idle and pays two CTA barriers and a warp-0 prefix for ~250 distance tests: 0.75 of the 0.82 ms rebuild at
// 1M beads went there (profiles/r1h_launches_host_kg1m.md).  Here every WARP owns a cell: its own staging slice, its own
// prefix over the <=27 adjacent cells, __syncwarp instead of __syncthreads, particles of the cell one after the other
// through the same nlist_scan_tile (same candidate order, same arithmetic => same list).  4 cells per CTA.
constexpr int NW_WARPS = 4;
constexpr int NW_CAP = 256; // staged candidates per warp and pass (multiple of 32)

template<bool MULTI_TYPE, bool HALF>
__global__ void __launch_bounds__(NW_WARPS * 32) k_nlist_warpcell(unsigned int* __restrict__ nlist,
                                                                  unsigned int* __restrict__ n_neigh,
                                                                  float4* __restrict__ last_pos,
                                                                  unsigned int* __restrict__ conditions,
                                                                  const unsigned int* __restrict__ Nmax,
                                                                  const unsigned int* __restrict__ head_list,
                                                                  unsigned int N,
                                                                  const unsigned int* __restrict__ cell_start,
                                                                  const float4* __restrict__ cell_xyzf,
                                                                  const unsigned int* __restrict__ cell_type, uint3 dim,
                                                                  Box box, const float* __restrict__ r_cut,
                                                                  const float* __restrict__ r_listsq, unsigned int ntypes,
                                                                  const unsigned int* __restrict__ enable, OverflowHook hook)
    {
    pdl_wait();
    if (enable && !*enable) return;
    extern __shared__ float4 s_all[];
    __shared__ unsigned int s_beg_all[NW_WARPS][28], s_off_all[NW_WARPS][28];
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float4* s_cand = s_all + warp * NW_CAP;
    unsigned int* s_type = reinterpret_cast<unsigned int*>(s_all + NW_WARPS * NW_CAP) + warp * NW_CAP;
    unsigned int* s_beg = s_beg_all[warp];
    unsigned int* s_off = s_off_all[warp];

    const unsigned int ncells = dim.x * dim.y * dim.z;
    const unsigned int cell = blockIdx.x * NW_WARPS + warp;
    if (cell >= ncells) return; // no CTA-wide barrier below: a warp may leave on its own
    const unsigned int own_beg = cell_start[cell], own_end = cell_start[cell + 1];
    if (own_beg == own_end) return;
    const unsigned int bx = cell % dim.x, by = (cell / dim.x) % dim.y, bz = cell / (dim.x * dim.y);

    int xs0, xs1, xs2, ys0, ys1, ys2, zs0, zs1, zs2;
    const int nx = axis_set3(bx, dim.x, box.periodic.x, xs0, xs1, xs2);
    const int ny = axis_set3(by, dim.y, box.periodic.y, ys0, ys1, ys2);
    const int nz = axis_set3(bz, dim.z, box.periodic.z, zs0, zs1, zs2);
    const unsigned int nr = (unsigned int)(nx * ny * nz);
    unsigned int beg = 0, len = 0;
    if (lane < nr)
        {
        // (c, q, r) in lexicographic order == ascending cell id (id = (z*dimy + y)*dimx + x)
        const int r = lane % nx, q = (lane / nx) % ny, c = lane / (nx * ny);
        const int x = (r == 0) ? xs0 : (r == 1) ? xs1 : xs2;
        const int y = (q == 0) ? ys0 : (q == 1) ? ys1 : ys2;
        const int z = (c == 0) ? zs0 : (c == 1) ? zs1 : zs2;
        const unsigned int nc = ((unsigned int)z * dim.y + (unsigned int)y) * dim.x + (unsigned int)x;
        beg = cell_start[nc];
        len = cell_start[nc + 1] - beg;
        }
    unsigned int incl = len;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
        {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((int)lane >= o) incl += t;
        }
    if (lane < 28)
        {
        s_beg[lane] = beg;
        s_off[lane] = incl - len;
        }
    const unsigned int total = __shfl_sync(0xffffffffu, incl, nr - 1);
    __syncwarp();
    const float rcut_00 = MULTI_TYPE ? 0.0f : r_cut[0];
    const float rlsq_00 = MULTI_TYPE ? 0.0f : r_listsq[0];
    const bool wx = cell_needs_image(box.iperiodic.x, box.periodic.x, box.iforce.x, box.ighost.x, dim.x, bx);
    const bool wy = cell_needs_image(box.iperiodic.y, box.periodic.y, box.iforce.y, box.ighost.y, dim.y, by);
    const bool wz = cell_needs_image(box.iperiodic.z, box.periodic.z, box.iforce.z, box.ighost.z, dim.z, bz);
    const bool wrap_any = wx || wy || wz;

    for (unsigned int pass = 0; pass < total; pass += NW_CAP)
        {
        const unsigned int cnt = min((unsigned int)NW_CAP, total - pass);
        const unsigned int cnt_padded = (cnt + 31u) & ~31u; // one candidate per lane and trip (U = 1): pad to 32 only
        if (pass) __syncwarp(); // the warp is done with the previous tile
        // lane r copies "its" adjacent cell (a handful of entries) into the staged order: no per-entry range search
        if (lane < nr)
            {
            const unsigned int off = incl - len; // == s_off[lane]
            const unsigned int k0 = pass > off ? pass - off : 0u;
            const unsigned int k1 = min(len, pass + NW_CAP > off ? pass + NW_CAP - off : 0u);
            for (unsigned int k = k0; k < k1; k++)
                {
                s_cand[off + k - pass] = cell_xyzf[beg + k];
                if (MULTI_TYPE) s_type[off + k - pass] = cell_type[beg + k];
                }
            }
        if (cnt + lane < cnt_padded)
            {
            // sentinel: dr^2 overflows to +inf, never <= r_list^2; index 0xffffffff is nobody
            s_cand[cnt + lane] = make_float4(1.0e30f, 1.0e30f, 1.0e30f, __int_as_float(-1));
            if (MULTI_TYPE) s_type[cnt + lane] = 0u;
            }
        __syncwarp();

        for (unsigned int p = own_beg; p < own_end; p++)
            {
            const float4 pi4 = cell_xyzf[p];
            const unsigned int i = (unsigned int)__float_as_int(pi4.w);
            if (i >= N) continue; // ghosts are candidates only
            const unsigned int type_i = MULTI_TYPE ? cell_type[p] : 0u;
            const unsigned int Nmax_i = Nmax[type_i];
            unsigned int* row = nlist + head_list[i];
            unsigned int n = pass ? n_neigh[i] : 0u;
            if (wrap_any)
                n = nlist_scan_tile<MULTI_TYPE, true, HALF, 1>(n, pi4, i, type_i, Nmax_i, row, s_cand, s_type, cnt_padded,
                                                            lane, box, wx, wy, wz, rcut_00, rlsq_00, r_cut, r_listsq,
                                                            ntypes);
            else
                n = nlist_scan_tile<MULTI_TYPE, false, HALF, 1>(n, pi4, i, type_i, Nmax_i, row, s_cand, s_type, cnt_padded,
                                                             lane, box, wx, wy, wz, rcut_00, rlsq_00, r_cut, r_listsq,
                                                             ntypes);
            if (lane == 0)
                {
                if (pass + NW_CAP < total)
                    n_neigh[i] = n;
                else
                    {
                    n_neigh[i] = min(n, Nmax_i);
                    if (n > Nmax_i) nlist_report_overflow(conditions, type_i, n, hook);
                    if (last_pos) last_pos[i] = make_float4(pi4.x, pi4.y, pi4.z, 0.0f);
                    }
                }
            __syncwarp();
            }
        }
    }

// ---- cell-scan build: one CTA per cell, EIGHT lanes per particle, each scanning a contiguous eighth of the staged
// candidates -------------------------------------------------------------------------------------------------------------
// The tiled kernel above spends one warp per particle with lanes = candidates: every 32 tests cost a ballot, two popc and
// a scattered store decision, ~1.9 warp-instructions per test, and the build is issue-bound (~100 % issue utilisation at
// 1M particles).  Here a warp holds 4 particles x 8 lanes; the candidate list is the same for every particle of the cell,
// so the scan loop is warp-uniform (no divergence, 8 distinct shared-memory addresses per load, the rest broadcast) and a
// test costs its arithmetic plus one bit of a register mask.  Lane q of a particle scans candidates [q*Q, (q+1)*Q); the
// eight sub-lists are concatenated (a width-8 prefix of the popcounts) into the row, which keeps the oracle's order
// (ascending staged index = neighbor cells by id, occupants by index).  The candidates are staged with one bulk copy
// (cp.async.bulk) per adjacent cell.  Membership arithmetic as above (bit-exact vs the oracle).
constexpr int NC_BLOCK = 128; // 4 warps = 16 particles per round
constexpr int NC_G = 8;       // lanes per particle
constexpr int NC_WORDS = 4;   // 32-candidate mask words per lane: Q <= 128, i.e. <= 1024 candidates per pass

template<bool MULTI_TYPE, bool WRAP, bool HALF>
__device__ __forceinline__ unsigned int nlist_scan_word(const float4 pi4, unsigned int i, unsigned int type_i,
                                                        const float4* s_cand, const unsigned int* s_type,
                                                        unsigned int t0, unsigned int nq, const Box& box, bool wx, bool wy,
                                                        bool wz, float rcut_00, float rlsq_00,
                                                        const float* __restrict__ r_cut,
                                                        const float* __restrict__ r_listsq, unsigned int ntypes)
    {
    unsigned int mask = 0, bit = 1;
#pragma unroll 2
    for (unsigned int q = 0; q < nq; q++)
        {
        const float4 cj = s_cand[t0 + q];
        const unsigned int j = (unsigned int)__float_as_int(cj.w);
        float rlsq = rlsq_00;
        bool excluded = (j == i);
        if (MULTI_TYPE)
            {
            const unsigned int tj = s_type[t0 + q];
            excluded |= __ldg(&r_cut[type_i * ntypes + tj]) <= 0.0f;
            rlsq = __ldg(&r_listsq[type_i * ntypes + tj]);
            }
        else
            excluded |= rcut_00 <= 0.0f;
        float3 dx = make_float3(__fsub_rn(pi4.x, cj.x), __fsub_rn(pi4.y, cj.y), __fsub_rn(pi4.z, cj.z));
        if (WRAP) dx = min_image_exact_sel(dx, box, wx, wy, wz);
        const bool hit = !excluded && (dot_exact(dx) <= rlsq) && (!HALF || i < j);
        if (hit) mask |= bit;
        bit += bit;
        }
    return mask;
    }

template<bool MULTI_TYPE, bool HALF, bool BULK>
__global__ void __launch_bounds__(NC_BLOCK) k_nlist_cellscan(unsigned int* __restrict__ nlist,
                                                             unsigned int* __restrict__ n_neigh,
                                                             float4* __restrict__ last_pos,
                                                             unsigned int* __restrict__ conditions,
                                                             const unsigned int* __restrict__ Nmax,
                                                             const unsigned int* __restrict__ head_list, unsigned int N,
                                                             const unsigned int* __restrict__ cell_start,
                                                             const float4* __restrict__ cell_xyzf,
                                                             const unsigned int* __restrict__ cell_type, uint3 dim,
                                                             Box box, const float* __restrict__ r_cut,
                                                             const float* __restrict__ r_listsq, unsigned int ntypes,
                                                             unsigned int cap, const unsigned int* __restrict__ enable, OverflowHook hook)
    {
    pdl_wait();
    if (enable && !*enable) return;
    // dynamic shared memory: [cap + 16] candidates, then [cap + 16] types
    extern __shared__ float4 s_cand[];
    unsigned int* s_type = reinterpret_cast<unsigned int*>(s_cand + cap + 16);
    __shared__ unsigned int s_beg[28], s_off[28];
    __shared__ unsigned int s_nr, s_heads;
    __shared__ __align__(8) unsigned long long s_bar;

    const unsigned int bx = blockIdx.x, by = blockIdx.y, bz = blockIdx.z;
    const unsigned int cell = (bz * dim.y + by) * dim.x + bx;
    const unsigned int own_beg = cell_start[cell], own_end = cell_start[cell + 1];
    if (own_beg == own_end) return; // uniform per CTA
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    if (warp == 0)
        {
        if (lane == 0 && BULK) mbar_init(&s_bar, 1);
        int xs[3], ys[3], zs[3];
        const int nx = axis_set(bx, dim.x, box.periodic.x, xs);
        const int ny = axis_set(by, dim.y, box.periodic.y, ys);
        const int nz = axis_set(bz, dim.z, box.periodic.z, zs);
        const int nr = nx * ny * nz;
        unsigned int beg = 0, len = 0, nc = 0xffffffffu;
        if ((int)lane < nr)
            {
            const int r = lane % nx, q = (lane / nx) % ny, c = lane / (nx * ny);
            const int x = (r == 0) ? xs[0] : (r == 1) ? xs[1] : xs[2];
            const int y = (q == 0) ? ys[0] : (q == 1) ? ys[1] : ys[2];
            const int z = (c == 0) ? zs[0] : (c == 1) ? zs[1] : zs[2];
            nc = ((unsigned int)z * dim.y + (unsigned int)y) * dim.x + (unsigned int)x;
            beg = cell_start[nc];
            len = cell_start[nc + 1] - beg;
            }
        unsigned int incl = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
            {
            const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
            if ((int)lane >= o) incl += t;
            }
        // cells with consecutive ids are consecutive in memory AND in the staged order: one bulk copy per run
        const unsigned int prev_nc = __shfl_up_sync(0xffffffffu, nc, 1);
        const bool head = (int)lane < nr && (lane == 0 || nc != prev_nc + 1u);
        const unsigned int heads = __ballot_sync(0xffffffffu, head);
        if (lane < 28)
            {
            s_beg[lane] = beg;
            s_off[lane] = incl - len;
            }
        if ((int)lane == nr - 1)
            {
            s_off[nr] = incl;
            s_nr = nr;
            s_heads = heads;
            }
        }
    __syncthreads();
    const unsigned int nr = s_nr;
    const unsigned int total = s_off[nr];
    const float rcut_00 = MULTI_TYPE ? 0.0f : r_cut[0];
    const float rlsq_00 = MULTI_TYPE ? 0.0f : r_listsq[0];
    const bool wx = cell_needs_image(box.iperiodic.x, box.periodic.x, box.iforce.x, box.ighost.x, dim.x, bx);
    const bool wy = cell_needs_image(box.iperiodic.y, box.periodic.y, box.iforce.y, box.ighost.y, dim.y, by);
    const bool wz = cell_needs_image(box.iperiodic.z, box.periodic.z, box.iforce.z, box.ighost.z, dim.z, bz);
    const bool wrap_any = wx || wy || wz;
    const unsigned int slot = threadIdx.x / NC_G, sub = threadIdx.x % NC_G;
    unsigned int parity = 0;

    for (unsigned int pass = 0; pass < total; pass += cap)
        {
        const unsigned int cnt = min(cap, total - pass);
        const unsigned int Q = (((cnt + NC_G - 1) / NC_G) + 1u) & ~1u; // per-lane share, even (the scan is unrolled by 2)
        const unsigned int cnt_padded = Q * NC_G;                      // <= cap + 15
        if (pass) __syncthreads(); // everyone is done with the previous window
        if (BULK)
            {
            // one bulk copy per run of adjacent cells (clipped to this pass's window), issued by the lanes of warp 0
            if (warp == 0)
                {
                if (pass) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                if (lane == 0) mbar_arrive_expect_tx(&s_bar, cnt * (unsigned int)sizeof(float4));
                __syncwarp();
                const unsigned int heads = s_heads;
                if ((heads >> lane) & 1u)
                    {
                    const unsigned int later = (lane < 31) ? (heads >> (lane + 1)) : 0u;
                    const unsigned int next = later ? lane + (unsigned int)__ffs(later) : nr; // next run's first cell
                    const unsigned int lo = max(s_off[lane], pass), hi = min(s_off[next], pass + cnt);
                    if (hi > lo)
                        bulk_g2s(s_cand + (lo - pass), cell_xyzf + s_beg[lane] + (lo - s_off[lane]),
                                 (hi - lo) * (unsigned int)sizeof(float4), &s_bar);
                    }
                }
            // sentinels: dr^2 overflows to +inf, never <= r_list^2; index 0xffffffff is nobody
            for (unsigned int t = cnt + threadIdx.x; t < cnt_padded; t += NC_BLOCK)
                s_cand[t] = make_float4(1.0e30f, 1.0e30f, 1.0e30f, __int_as_float(-1));
            __syncthreads();
            mbar_wait(&s_bar, parity);
            parity ^= 1u;
            }
        else
            {
            for (unsigned int t = threadIdx.x; t < cnt_padded; t += NC_BLOCK)
                {
                if (t < cnt)
                    {
                    const unsigned int g = pass + t;
                    unsigned int lo = 0, hi = nr; // adjacent cell holding staged entry g: last r with s_off[r] <= g
                    while (hi - lo > 1)
                        {
                        const unsigned int mid = (lo + hi) >> 1;
                        if (s_off[mid] <= g) lo = mid; else hi = mid;
                        }
                    const unsigned int e = s_beg[lo] + (g - s_off[lo]);
                    s_cand[t] = cell_xyzf[e];
                    if (MULTI_TYPE) s_type[t] = cell_type[e];
                    }
                else
                    {
                    s_cand[t] = make_float4(1.0e30f, 1.0e30f, 1.0e30f, __int_as_float(-1));
                    if (MULTI_TYPE) s_type[t] = 0u;
                    }
                }
            __syncthreads();
            }

        for (unsigned int p0 = own_beg; p0 < own_end; p0 += NC_BLOCK / NC_G)
            {
            // a warp whose 4 particle slots are all past the end has nothing to do (warp-uniform)
            if (p0 + warp * (32 / NC_G) >= own_end) continue;
            const unsigned int p = p0 + slot;
            const bool in_cell = p < own_end;
            const float4 pi4 = in_cell ? cell_xyzf[p] : make_float4(0.0f, 0.0f, 0.0f, __int_as_float(-2));
            const unsigned int i = (unsigned int)__float_as_int(pi4.w);
            const bool active = in_cell && i < N; // ghosts are candidates only
            const unsigned int type_i = (MULTI_TYPE && active) ? cell_type[p] : 0u;
            const unsigned int t0 = sub * Q;
            unsigned int masks[NC_WORDS];
            unsigned int c = 0;
#pragma unroll
            for (int w = 0; w < NC_WORDS; w++)
                {
                masks[w] = 0u;
                if (w * 32u < Q)
                    {
                    const unsigned int nq = min(32u, Q - w * 32u);
                    if (wrap_any)
                        masks[w] = nlist_scan_word<MULTI_TYPE, true, HALF>(pi4, i, type_i, s_cand, s_type, t0 + w * 32u, nq,
                                                                           box, wx, wy, wz, rcut_00, rlsq_00, r_cut,
                                                                           r_listsq, ntypes);
                    else
                        masks[w] = nlist_scan_word<MULTI_TYPE, false, HALF>(pi4, i, type_i, s_cand, s_type, t0 + w * 32u, nq,
                                                                            box, wx, wy, wz, rcut_00, rlsq_00, r_cut,
                                                                            r_listsq, ntypes);
                    if (!active) masks[w] = 0u;
                    c += __popc(masks[w]);
                    }
                }
            // concatenate the eight sub-lists of a particle: exclusive prefix over its lanes
            unsigned int incl = c;
#pragma unroll
            for (int o = 1; o < NC_G; o <<= 1)
                {
                const unsigned int v = __shfl_up_sync(0xffffffffu, incl, o, NC_G);
                if (sub >= (unsigned int)o) incl += v;
                }
            const unsigned int found = __shfl_sync(0xffffffffu, incl, NC_G - 1, NC_G);
            const unsigned int Nmax_i = active ? Nmax[type_i] : 0u;
            const unsigned int n0 = (pass && active) ? n_neigh[i] : 0u; // running (unclamped) count of earlier passes
            if (active)
                {
                unsigned int* row = nlist + head_list[i];
                unsigned int k = n0 + incl - c;
#pragma unroll
                for (int w = 0; w < NC_WORDS; w++)
                    {
                    unsigned int m = masks[w];
                    while (m)
                        {
                        const unsigned int q = __ffs(m) - 1u;
                        m &= m - 1u;
                        if (k < Nmax_i) row[k] = (unsigned int)__float_as_int(s_cand[t0 + w * 32u + q].w);
                        k++;
                        }
                    }
                }
            __syncwarp();
            if (active && sub == 0)
                {
                const unsigned int n = n0 + found;
                if (pass + cap < total)
                    n_neigh[i] = n;
                else
                    {
                    n_neigh[i] = min(n, Nmax_i);
                    if (n > Nmax_i) nlist_report_overflow(conditions, type_i, n, hook);
                    if (last_pos) last_pos[i] = make_float4(pi4.x, pi4.y, pi4.z, 0.0f);
                    }
                }
            }
        }
    }

// ---- distance check (NeighborList::distanceCheck) -----------------------------------------------------------------
__global__ void __launch_bounds__(256) k_nlist_dist_check(unsigned int* __restrict__ result,
                                                          const float4* __restrict__ last_pos,
                                                          const float4* __restrict__ pos, unsigned int N, Box box,
                                                          const float* __restrict__ rcut_max, float r_buff,
                                                          float lambda_min, float3 lambda, unsigned int checkn)
    {
    unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool moved = false;
    if (i < N)
        {
        const float4 p = pos[i];
        const float4 lp = last_pos[i];
        const float old_rmin = rcut_max[(unsigned int)__float_as_int(p.w)];
        const float rmax = __fadd_rn(old_rmin, r_buff);
        const float delta_max = __fdiv_rn(__fsub_rn(__fmul_rn(rmax, lambda_min), old_rmin), 2.0f);
        const float maxsq = (delta_max > 0.0f) ? __fmul_rn(delta_max, delta_max) : 0.0f;
        float3 dx = make_float3(__fsub_rn(p.x, __fmul_rn(lambda.x, lp.x)), __fsub_rn(p.y, __fmul_rn(lambda.y, lp.y)),
                                __fsub_rn(p.z, __fmul_rn(lambda.z, lp.z)));
        dx = min_image_exact(dx, box);
        moved = dot_exact(dx) >= maxsq;
        }
    // one write per warp that saw a mover (all writers store the same value)
    if (__any_sync(0xffffffffu, moved) && (threadIdx.x & 31) == 0) atomicMax(result, checkn);
    }

// ---- head list: exclusive scan of Nmax[type_i] -----------------------------------------------------------------------
// Tile = 256 threads x 8 particles.  reduce -> single-CTA scan of tile sums -> apply.
constexpr int HL_ITEMS = 8;
constexpr int HL_TILE = 256 * HL_ITEMS;

__device__ __forceinline__ unsigned int block_exclusive_scan_256(unsigned int v, unsigned int* total)
    {
    __shared__ unsigned int ws[8];
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
        {
        unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
        }
    if (lane == 31) ws[warp] = incl;
    __syncthreads();
    unsigned int off = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < 8; w++)
        {
        unsigned int s = ws[w];
        if (w < (int)warp) off += s;
        tot += s;
        }
    __syncthreads();
    *total = tot;
    return off + incl - v;
    }

__global__ void __launch_bounds__(256) k_head_reduce(unsigned int* __restrict__ tile_sums,
                                                     const unsigned int* __restrict__ Nmax,
                                                     const float4* __restrict__ pos, unsigned int N)
    {
    unsigned int base = blockIdx.x * HL_TILE + threadIdx.x * HL_ITEMS;
    unsigned int s = 0;
#pragma unroll
    for (int k = 0; k < HL_ITEMS; k++)
        if (base + k < N) s += Nmax[(unsigned int)__float_as_int(pos[base + k].w)];
    unsigned int tot;
    block_exclusive_scan_256(s, &tot);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
    }

__global__ void __launch_bounds__(256) k_head_scan_tiles(unsigned int* __restrict__ tile_sums, unsigned int ntiles,
                                                         unsigned int* __restrict__ req_size)
    {
    unsigned int carry = 0;
    for (unsigned int base = 0; base < ntiles; base += 256)
        {
        unsigned int v = (base + threadIdx.x < ntiles) ? tile_sums[base + threadIdx.x] : 0u;
        unsigned int tot;
        unsigned int ex = block_exclusive_scan_256(v, &tot);
        if (base + threadIdx.x < ntiles) tile_sums[base + threadIdx.x] = carry + ex;
        carry += tot;
        }
    if (threadIdx.x == 0 && req_size) *req_size = carry;
    }

__global__ void __launch_bounds__(256) k_head_apply(unsigned int* __restrict__ head_list,
                                                    const unsigned int* __restrict__ tile_sums,
                                                    const unsigned int* __restrict__ Nmax,
                                                    const float4* __restrict__ pos, unsigned int N)
    {
    unsigned int base = blockIdx.x * HL_TILE + threadIdx.x * HL_ITEMS;
    unsigned int v[HL_ITEMS];
    unsigned int s = 0;
#pragma unroll
    for (int k = 0; k < HL_ITEMS; k++)
        {
        v[k] = (base + k < N) ? Nmax[(unsigned int)__float_as_int(pos[base + k].w)] : 0u;
        s += v[k];
        }
    unsigned int tot;
    unsigned int ex = block_exclusive_scan_256(s, &tot) + tile_sums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < HL_ITEMS; k++)
        {
        if (base + k < N) head_list[base + k] = ex;
        ex += v[k];
        }
    }

// single-type closed form: head[i] = i * Nmax[0]
__global__ void __launch_bounds__(256) k_head_uniform(unsigned int* __restrict__ head_list,
                                                      unsigned int* __restrict__ req_size,
                                                      const unsigned int* __restrict__ Nmax, unsigned int N)
    {
    unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int m = Nmax[0];
    if (i < N) head_list[i] = i * m;
    if (i == 0 && req_size) *req_size = N * m;
    }
} // namespace b200

using namespace b200;

extern "C" {

int b200_nlist_build_head_list(unsigned int* d_head_list, unsigned int* d_req_size_nlist, const unsigned int* d_Nmax,
                               const float* d_pos, unsigned int N, unsigned int ntypes, b200_stream_t stream_)
    {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (N == 0)
        {
        if (d_req_size_nlist) cudaMemsetAsync(d_req_size_nlist, 0, sizeof(unsigned int), stream);
        return (int)cudaGetLastError();
        }
    if (ntypes == 1)
        {
        k_head_uniform<<<(N + 255) / 256, 256, 0, stream>>>(d_head_list, d_req_size_nlist, d_Nmax, N);
        return (int)cudaGetLastError();
        }
    // the tile sums live at the tail of the head list's own allocation?  No: callers own memory; use a small
    // stream-ordered allocation (ntiles * 4 B, e.g. 16 kB at 8M particles).
    const unsigned int ntiles = (N + HL_TILE - 1) / HL_TILE;
    unsigned int* d_tiles = nullptr;
    cudaError_t err = cudaMallocAsync((void**)&d_tiles, sizeof(unsigned int) * ntiles, stream);
    if (err != cudaSuccess) return (int)err;
    k_head_reduce<<<ntiles, 256, 0, stream>>>(d_tiles, d_Nmax, (const float4*)d_pos, N);
    k_head_scan_tiles<<<1, 256, 0, stream>>>(d_tiles, ntiles, d_req_size_nlist);
    k_head_apply<<<ntiles, 256, 0, stream>>>(d_head_list, d_tiles, d_Nmax, (const float4*)d_pos, N);
    cudaFreeAsync(d_tiles, stream);
    return (int)cudaGetLastError();
    }

} // extern "C" (reopened below)

namespace b200
{
cudaError_t launch_nlist_binned(const b200_nlist_args* a, const unsigned int* enable, cudaStream_t stream,
                                OverflowHook hook)
    {
    if (!box_is_orthorhombic(a->box)) return cudaErrorNotSupported;
    if (a->N == 0) return cudaSuccess;
    if (a->ntypes > 1 && !a->d_cell_type) return cudaErrorInvalidValue;
    Box box = make_box_image(a->box, a->image_box);
    uint3 dim = make_uint3(a->dim.x, a->dim.y, a->dim.z);
    float3 gw = make_float3(a->ghost_width[0], a->ghost_width[1], a->ghost_width[2]);
        {
        // ghost cells on a face: those below ghost_width / cell_width (bins span [lo - gw, hi + gw)); one more layer has them
        // among its neighbors
        unsigned char ig[3];
        const unsigned int d3[3] = {a->dim.x, a->dim.y, a->dim.z};
        for (int ax = 0; ax < 3; ax++)
            {
            const float cw = (a->box.L[ax] + 2.0f * a->ghost_width[ax]) / (float)d3[ax];
            const float layers = a->ghost_width[ax] > 0.0f ? a->ghost_width[ax] / cw + 1e-3f : 0.0f;
            ig[ax] = (unsigned char)std::min(254.0f, std::floor(layers) + 1.0f);
            }
        box.ighost = make_uchar3(ig[0], ig[1], ig[2]);
        }
    const bool pdl = enable && pdl_enabled(); // chained behind k_cell_rank (launch_chain, common.cuh)
    static const bool use_warp_kernel = getenv("HOOMD_B200_NLIST_WARP") != nullptr; // A/B switches for profiling
    // the cell-scan kernel executes ~2.5x fewer instructions than the tiled one but is latency-bound per CTA and measured
    // slower on B200 (rebuild 85 vs 71 us at 64k particles, 782 vs 637 us at 1M): opt-in until that is fixed
    static const bool use_cellscan_kernel = getenv("HOOMD_B200_NLIST_CELLSCAN") != nullptr;
    if (!use_warp_kernel && use_cellscan_kernel)
        {
        if (a->dim.y > 65535u || a->dim.z > 65535u) return cudaErrorInvalidValue;
        const dim3 grid(a->dim.x, a->dim.y, a->dim.z);
        // candidates staged per pass: the caller's estimate of a cell neighborhood's population (cand_hint), else 512
        unsigned int cap = a->cand_hint ? a->cand_hint : 512u;
        cap = ((cap < 64u ? 64u : cap > 1024u ? 1024u : cap) + 7u) & ~7u;
        size_t smem = (cap + 16) * sizeof(float4);
        if (a->ntypes > 1) smem += (cap + 16) * sizeof(unsigned int);
        static const bool manual_staging = getenv("HOOMD_B200_NC_MANUAL") != nullptr;
#define B200_NC_LAUNCH(MT, HALF) do { if (!MT && !manual_staging) B200_NC_LAUNCH2(MT, HALF, !MT); else B200_NC_LAUNCH2(MT, HALF, false); } while (0)
#define B200_NC_LAUNCH2(MT, HALF, BULK)                                                                              \
    do {                                                                                                             \
        cudaFuncSetAttribute(k_nlist_cellscan<MT, HALF, BULK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        k_nlist_cellscan<MT, HALF, BULK><<<grid, NC_BLOCK, smem, stream>>>(                                          \
            a->d_nlist, a->d_n_neigh, (float4*)a->d_last_pos, a->d_conditions, a->d_Nmax, a->d_head_list, a->N,      \
            a->d_cell_start, (const float4*)a->d_cell_xyzf, a->d_cell_type, dim, box, a->d_r_cut, a->d_r_listsq,     \
            a->ntypes, cap, enable, hook);                                                                               \
    } while (0)
        if (a->ntypes > 1)
            {
            if (a->half_storage) B200_NC_LAUNCH(true, true); else B200_NC_LAUNCH(true, false);
            }
        else
            {
            if (a->half_storage) B200_NC_LAUNCH(false, true); else B200_NC_LAUNCH(false, false);
            }
#undef B200_NC_LAUNCH
#undef B200_NC_LAUNCH2
        return cudaGetLastError();
        }
    // sparse cells (<= ~5 particles per cell): one warp per cell instead of one CTA (see k_nlist_warpcell)
    static const char* warpcell_env = getenv("HOOMD_B200_NLIST_WARPCELL");
    const bool sparse = a->cand_hint && a->cand_hint <= 250u;
    if (!use_warp_kernel && !use_cellscan_kernel && (warpcell_env ? atoi(warpcell_env) != 0 : sparse))
        {
        const unsigned int ncells = a->dim.x * a->dim.y * a->dim.z;
        const unsigned int grid = (ncells + NW_WARPS - 1) / NW_WARPS;
        const size_t smem = (size_t)NW_WARPS * NW_CAP * (sizeof(float4) + (a->ntypes > 1 ? sizeof(unsigned int) : 0));
#define B200_NW_LAUNCH(MT, HALF)                                                                                     \
    launch_chain(k_nlist_warpcell<MT, HALF>, dim3(grid), dim3(NW_WARPS * 32), smem, stream, pdl,                  \
        a->d_nlist, a->d_n_neigh, (float4*)a->d_last_pos, a->d_conditions, a->d_Nmax, a->d_head_list, a->N,         \
        a->d_cell_start, (const float4*)a->d_cell_xyzf, a->d_cell_type, dim, box, a->d_r_cut, a->d_r_listsq,        \
        a->ntypes, enable, hook)
        if (a->ntypes > 1)
            {
            if (a->half_storage) B200_NW_LAUNCH(true, true); else B200_NW_LAUNCH(true, false);
            }
        else
            {
            if (a->half_storage) B200_NW_LAUNCH(false, true); else B200_NW_LAUNCH(false, false);
            }
#undef B200_NW_LAUNCH
        return cudaGetLastError();
        }
    if (!use_warp_kernel)
        {
        if (a->dim.y > 65535u || a->dim.z > 65535u) return cudaErrorInvalidValue;
        const dim3 grid(a->dim.x, a->dim.y, a->dim.z);
        const size_t smem1 = NL_CAP * sizeof(float4), smemN = NL_CAP * (sizeof(float4) + sizeof(unsigned int));
#define B200_NL_LAUNCH(MT, HALF, SMEM)                                                                               \
    launch_chain(k_nlist_tiled<MT, HALF>, grid, dim3(NL_BLOCK), SMEM, stream, pdl,                                 \
        a->d_nlist, a->d_n_neigh, (float4*)a->d_last_pos, a->d_conditions, a->d_Nmax, a->d_head_list, a->N,         \
        a->d_cell_start, (const float4*)a->d_cell_xyzf, a->d_cell_type, dim, box, a->d_r_cut, a->d_r_listsq,        \
        a->ntypes, enable, hook)
        if (a->ntypes > 1)
            {
            if (a->half_storage) B200_NL_LAUNCH(true, true, smemN); else B200_NL_LAUNCH(true, false, smemN);
            }
        else
            {
            if (a->half_storage) B200_NL_LAUNCH(false, true, smem1); else B200_NL_LAUNCH(false, false, smem1);
            }
#undef B200_NL_LAUNCH
        return cudaGetLastError();
        }
    const unsigned int warps_per_block = 4;
    const unsigned int grid = (a->N + warps_per_block - 1) / warps_per_block;
    if (a->ntypes > 1)
        k_nlist_binned<true><<<grid, 128, 0, stream>>>(a->d_nlist, a->d_n_neigh, (float4*)a->d_last_pos, a->d_conditions,
                                                       a->d_Nmax, a->d_head_list, (const float4*)a->d_pos, a->N,
                                                       a->d_cell_start, (const float4*)a->d_cell_xyzf, a->d_cell_type,
                                                       dim, box, gw, a->d_r_cut, a->d_r_listsq, a->ntypes,
                                                       a->half_storage, enable, hook);
    else
        k_nlist_binned<false><<<grid, 128, 0, stream>>>(a->d_nlist, a->d_n_neigh, (float4*)a->d_last_pos,
                                                        a->d_conditions, a->d_Nmax, a->d_head_list,
                                                        (const float4*)a->d_pos, a->N, a->d_cell_start,
                                                        (const float4*)a->d_cell_xyzf, a->d_cell_type, dim, box, gw,
                                                        a->d_r_cut, a->d_r_listsq, a->ntypes, a->half_storage,
                                                        enable, hook);
    return cudaGetLastError();
    }
} // namespace b200

extern "C" {

int b200_compute_nlist_binned(const b200_nlist_args* a, b200_stream_t stream)
    {
    return (int)b200::launch_nlist_binned(a, nullptr, (cudaStream_t)stream);
    }

int b200_nlist_needs_update_check(unsigned int* d_result, const float* d_last_pos, const float* d_pos, unsigned int N,
                                  const b200_box* box, const float* d_rcut_max, float r_buff, unsigned int ntypes,
                                  float lambda_min, const float lambda[3], unsigned int checkn, b200_stream_t stream)
    {
    (void)ntypes;
    if (!box_is_orthorhombic(*box)) return (int)cudaErrorNotSupported;
    if (N == 0) return 0;
    k_nlist_dist_check<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        d_result, (const float4*)d_last_pos, (const float4*)d_pos, N, make_box(*box), d_rcut_max, r_buff, lambda_min,
        make_float3(lambda[0], lambda[1], lambda[2]), checkn);
    return (int)cudaGetLastError();
    }

} // extern "C"
