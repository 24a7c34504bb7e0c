// CellList for sm_100a: count -> scan -> scatter -> in-cell rank.  Replaces hoomd/CellListGPU.cu:40-219 (atomicInc into a
// padded Ncells x Nmax array) with contiguous, index-ordered cell ranges.  See include/hoomd_b200.h for the layout.
#include "common.cuh"
#include <cooperative_groups.h>
#include <cstdlib>

namespace b200
{
namespace cg = cooperative_groups;
// ---- K1: bin every particle, count cell occupancies ---------------------------------------------------------------
// 1 thread / particle, coalesced float4 load.  atomicAdd on cell_count returns an arbitrary slot; the final in-cell
// order is fixed later by rank (k_cell_rank), so the result is deterministic.
__global__ void __launch_bounds__(256) k_cell_bin(unsigned int* __restrict__ cell_count,
                                                  unsigned int* __restrict__ particle_cell,
                                                  unsigned int* __restrict__ particle_slot,
                                                  unsigned int* __restrict__ conditions,
                                                  const float4* __restrict__ pos, unsigned int N, unsigned int n_ghost,
                                                  Box box, uint3 dim, float3 ghost_width,
                                                  const unsigned int* __restrict__ enable,
                                                  const unsigned int* __restrict__ d_N,
                                                  const unsigned char* __restrict__ ghost_dir,
                                                  const float* __restrict__ ghost_shift)
    {
    pdl_trigger(); // first kernel of the rebuild chain: let the scan's CTAs be scheduled behind this grid
    if (enable && !*enable) return; // device-driven stepping: the rebuild was not requested this step
    unsigned int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N + n_ghost) return;
    if (d_N)
        {
        // multi-GPU layout: locals [0, *d_N), unused capacity [*d_N, N), ghosts [N, N + n_ghost)
        const unsigned int Nloc = *d_N;
        if (n >= Nloc && n < N)
            {
            particle_cell[n] = B200_NOT_BINNED;
            return;
            }
        }
    float4 p = pos[n];
    if (ghost_dir && n >= N)
        {
        // a ghost with raw coordinates is binned at its periodic image next to the local box
        const float* sh = ghost_shift + 3u * ghost_dir[n - N];
        p.x = __fadd_rn(p.x, sh[0]);
        p.y = __fadd_rn(p.y, sh[1]);
        p.z = __fadd_rn(p.z, sh[2]);
        }
    unsigned int cell = B200_NOT_BINNED, slot = 0;
    if (isnan(p.x) || isnan(p.y) || isnan(p.z))
        atomicMax(&conditions[1], n + 1);
    else
        {
        int3 b;
        if (particle_bin(make_float3(p.x, p.y, p.z), box, ghost_width, dim, b))
            {
            cell = ((unsigned int)b.z * dim.y + (unsigned int)b.y) * dim.x + (unsigned int)b.x;
            slot = atomicAdd(&cell_count[cell], 1u);
            }
        else if (n < N)
            atomicMax(&conditions[2], n + 1); // ghosts out of range are silently ignored (CellList.cc:494-498)
        }
    particle_cell[n] = cell;
    particle_slot[n] = slot;
    }

// ---- K2: exclusive scan of the occupancies (single CTA, 1024 threads, warp-shuffle scan, sequential tiles) --------
// ncells is 8e3 (64k particles) .. 1e6 (8M particles); the scan reads 4 B/cell once and is far off the critical path.
__global__ void __launch_bounds__(1024) k_cell_scan(unsigned int* __restrict__ cell_start,
                                                    unsigned int* __restrict__ cell_count, unsigned int ncells,
                                                    unsigned int* __restrict__ conditions,
                                                    const unsigned int* __restrict__ enable)
    {
    pdl_wait();
    pdl_trigger();
    if (enable && !*enable) return;
    __shared__ unsigned int warp_sums[32];
    __shared__ unsigned int carry_s;
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    unsigned int max_occ = 0;
    __syncthreads();
    // each thread owns 4 consecutive cells per tile (16 B vector load when aligned)
    for (unsigned int base = 0; base < ncells; base += 4096)
        {
        unsigned int i0 = base + threadIdx.x * 4;
        const unsigned int carry = carry_s; // read before the barrier below; warp 0 rewrites it after that barrier
        unsigned int c[4];
#pragma unroll
        for (int k = 0; k < 4; k++)
            {
            c[k] = (i0 + k < ncells) ? cell_count[i0 + k] : 0u;
            // self-cleaning: the counts are consumed here, leave zeros behind for the next build (no memset node needed)
            if (i0 + k < ncells) cell_count[i0 + k] = 0u;
            }
        max_occ = max(max(max_occ, max(c[0], c[1])), max(c[2], c[3]));
        unsigned int tsum = c[0] + c[1] + c[2] + c[3];
        unsigned int incl = tsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
            {
            unsigned int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
            }
        if (lane == 31) warp_sums[warp] = incl;
        __syncthreads();
        if (warp == 0)
            {
            unsigned int w = warp_sums[lane];
            unsigned int wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1)
                {
                unsigned int v = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += v;
                }
            warp_sums[lane] = wi - w; // exclusive
            if (lane == 31) carry_s = carry + wi;
            }
        __syncthreads();
        unsigned int excl = carry + warp_sums[warp] + incl - tsum;
#pragma unroll
        for (int k = 0; k < 4; k++)
            {
            if (i0 + k < ncells) cell_start[i0 + k] = excl;
            excl += c[k];
            }
        __syncthreads();
        }
    if (threadIdx.x == 0) cell_start[ncells] = carry_s;
    // conditions.x = largest occupancy (what the reference's overflow protocol would grow Nmax to)
    for (int o = 16; o > 0; o >>= 1) max_occ = max(max_occ, __shfl_xor_sync(0xffffffffu, max_occ, o));
    if (lane == 0 && max_occ) atomicMax(&conditions[0], max_occ);
    }

// ---- K2 for large grids: the same scan over many CTAs in two launches ------------------------------------------------------
// The single CTA above walks ncells/4096 tiles one after the other: 66 us at 125k cells (LJ 1M) and 171 us at 328k cells
// (Kremer-Grest 1M, r_list 1.52) -- 10-14 % of a list rebuild (profiles/r1h_launches_host_*.md).  Here every CTA scans one
// 4096-cell tile (local exclusive scan + tile total), then a second launch adds the sum of the preceding tile totals
// (<= a few hundred values, re-summed by every CTA: cheaper than a third launch or a look-back chain).
constexpr unsigned int SCAN_TILE = 4096;

__device__ __forceinline__ unsigned int block_scan_1024(unsigned int tsum, unsigned int* warp_sums, unsigned int& total)
    {
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned int incl = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
        {
        unsigned int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
        }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0)
        {
        unsigned int w = warp_sums[lane];
        unsigned int wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
            {
            unsigned int v = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += v;
            }
        warp_sums[lane] = wi - w; // exclusive
        if (lane == 31) warp_sums[32] = wi;
        }
    __syncthreads();
    total = warp_sums[32];
    return warp_sums[warp] + incl - tsum; // exclusive prefix of this thread
    }

__global__ void __launch_bounds__(1024) k_cell_scan_tiles(unsigned int* __restrict__ cell_start,
                                                          unsigned int* __restrict__ cell_count, unsigned int ncells,
                                                          unsigned int* __restrict__ tile_sums,
                                                          unsigned int* __restrict__ conditions,
                                                          const unsigned int* __restrict__ enable)
    {
    pdl_wait();
    pdl_trigger();
    if (enable && !*enable) return;
    __shared__ unsigned int warp_sums[33];
    const unsigned int i0 = blockIdx.x * SCAN_TILE + threadIdx.x * 4;
    unsigned int c[4];
    if (i0 + 3 < ncells)
        {
        const uint4 v = *reinterpret_cast<const uint4*>(cell_count + i0); // cudaMalloc base + multiple of 16 B
        c[0] = v.x; c[1] = v.y; c[2] = v.z; c[3] = v.w;
        *reinterpret_cast<uint4*>(cell_count + i0) = make_uint4(0u, 0u, 0u, 0u); // self-cleaning, as above
        }
    else
        {
#pragma unroll
        for (int k = 0; k < 4; k++)
            {
            c[k] = (i0 + k < ncells) ? cell_count[i0 + k] : 0u;
            if (i0 + k < ncells) cell_count[i0 + k] = 0u;
            }
        }
    unsigned int max_occ = max(max(c[0], c[1]), max(c[2], c[3]));
    unsigned int total;
    unsigned int excl = block_scan_1024(c[0] + c[1] + c[2] + c[3], warp_sums, total);
#pragma unroll
    for (int k = 0; k < 4; k++)
        {
        if (i0 + k < ncells) cell_start[i0 + k] = excl;
        excl += c[k];
        }
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
    for (int o = 16; o > 0; o >>= 1) max_occ = max(max_occ, __shfl_xor_sync(0xffffffffu, max_occ, o));
    if ((threadIdx.x & 31) == 0 && max_occ) atomicMax(&conditions[0], max_occ);
    }

__global__ void __launch_bounds__(1024) k_cell_scan_fix(unsigned int* __restrict__ cell_start, unsigned int ncells,
                                                        const unsigned int* __restrict__ tile_sums,
                                                        const unsigned int* __restrict__ enable)
    {
    pdl_wait();
    pdl_trigger();
    if (enable && !*enable) return;
    __shared__ unsigned int warp_sums[33];
    // offset of this tile = sum of the totals of the tiles before it
    unsigned int part = 0;
    for (unsigned int t = threadIdx.x; t < blockIdx.x; t += 1024) part += tile_sums[t];
    unsigned int offset;
    block_scan_1024(part, warp_sums, offset);
    const unsigned int i0 = blockIdx.x * SCAN_TILE + threadIdx.x * 4;
    if (blockIdx.x)
        {
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (i0 + k < ncells) cell_start[i0 + k] += offset;
        }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) cell_start[ncells] = offset + tile_sums[blockIdx.x];
    }

// ---- K3: scatter particle indices to their (arbitrary) slot inside the cell range ----------------------------------
__global__ void __launch_bounds__(256) k_cell_scatter(unsigned int* __restrict__ scratch_idx,
                                                      const unsigned int* __restrict__ cell_start,
                                                      const unsigned int* __restrict__ particle_cell,
                                                      const unsigned int* __restrict__ particle_slot, unsigned int ntot,
                                                      const unsigned int* __restrict__ enable)
    {
    pdl_wait();
    pdl_trigger();
    if (enable && !*enable) return;
    unsigned int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= ntot) return;
    unsigned int cell = particle_cell[n];
    if (cell == B200_NOT_BINNED) return;
    scratch_idx[cell_start[cell] + particle_slot[n]] = n;
    }

// ---- K2 + K3 in one launch for small grids -------------------------------------------------------------------------------------
// At 64k particles (8000 cells) the single-CTA scan runs 3 us and costs one more dependent launch inside the rebuild body.
// Here EVERY scatter CTA scans all occupancies itself (ncells <= SS_MAX_CELLS: <= 32 kB of L2-resident counts per CTA, one
// 512-thread block scan, result in shared memory) and scatters its particles from that; CTA 0 also publishes cell_start and
// the largest occupancy.  The counts cannot be cleaned here (other CTAs still read them): k_cell_rank does it.
constexpr unsigned int SS_BLOCK = 512, SS_CELLS_PER = 16, SS_MAX_CELLS = SS_BLOCK * SS_CELLS_PER; // 8192 cells: 64k particles at rho 0.2

__global__ void __launch_bounds__(SS_BLOCK) k_cell_scan_scatter(unsigned int* __restrict__ cell_start,
                                                                const unsigned int* __restrict__ cell_count,
                                                                unsigned int ncells, unsigned int* __restrict__ conditions,
                                                                unsigned int* __restrict__ scratch_idx,
                                                                const unsigned int* __restrict__ particle_cell,
                                                                const unsigned int* __restrict__ particle_slot,
                                                                unsigned int ntot, const unsigned int* __restrict__ enable)
    {
    pdl_wait();
    pdl_trigger();
    if (enable && !*enable) return;
    __shared__ unsigned int s_start[SS_MAX_CELLS + 1];
    __shared__ unsigned int warp_sums[SS_BLOCK / 32 + 1];
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // this thread's particle: requested before the scan so that the loads overlap it
    const unsigned int n = blockIdx.x * SS_BLOCK + threadIdx.x;
    unsigned int cell = B200_NOT_BINNED, slot = 0;
    if (n < ntot)
        {
        cell = particle_cell[n];
        slot = particle_slot[n];
        }
    const unsigned int i0 = threadIdx.x * SS_CELLS_PER;
    unsigned int c[SS_CELLS_PER];
    unsigned int tsum = 0, max_occ = 0;
#pragma unroll
    for (unsigned int k = 0; k < SS_CELLS_PER; k += 4)
        {
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (i0 + k + 3 < ncells)
            v = *reinterpret_cast<const uint4*>(cell_count + i0 + k); // cudaMalloc base + multiple of 16 B
        else
            {
            if (i0 + k < ncells) v.x = cell_count[i0 + k];
            if (i0 + k + 1 < ncells) v.y = cell_count[i0 + k + 1];
            if (i0 + k + 2 < ncells) v.z = cell_count[i0 + k + 2];
            }
        c[k] = v.x; c[k + 1] = v.y; c[k + 2] = v.z; c[k + 3] = v.w;
        tsum += v.x + v.y + v.z + v.w;
        max_occ = max(max(max_occ, max(v.x, v.y)), max(v.z, v.w));
        }
    unsigned int incl = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
        {
        const unsigned int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
        }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0)
        {
        const unsigned int w = lane < SS_BLOCK / 32 ? warp_sums[lane] : 0u;
        unsigned int wi = w;
#pragma unroll
        for (int o = 1; o < SS_BLOCK / 32; o <<= 1)
            {
            const unsigned int v = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += v;
            }
        if (lane < SS_BLOCK / 32) warp_sums[lane] = wi - w; // exclusive
        if (lane == SS_BLOCK / 32 - 1) warp_sums[SS_BLOCK / 32] = wi;
        }
    __syncthreads();
    unsigned int excl = warp_sums[warp] + incl - tsum;
    const bool publish = blockIdx.x == 0;
#pragma unroll
    for (unsigned int k = 0; k < SS_CELLS_PER; k++)
        {
        if (i0 + k < ncells)
            {
            s_start[i0 + k] = excl;
            if (publish) cell_start[i0 + k] = excl;
            }
        excl += c[k];
        }
    if (publish)
        {
        if (threadIdx.x == 0) cell_start[ncells] = warp_sums[SS_BLOCK / 32];
        for (int o = 16; o > 0; o >>= 1) max_occ = max(max_occ, __shfl_xor_sync(0xffffffffu, max_occ, o));
        if (lane == 0 && max_occ) atomicMax(&conditions[0], max_occ);
        }
    __syncthreads();
    if (cell != B200_NOT_BINNED) scratch_idx[s_start[cell] + slot] = n;
    }

// ---- K4: rank every entry inside its cell by particle index and emit the final cell-sorted arrays -------------------
// 1 thread / entry.  A cell holds ~8 (rho=0.2) .. ~30 (rho=1.2) entries, so the rank loop is a handful of L1-resident
// 4 B loads; output writes are 16 B and land inside one contiguous cell range.
__global__ void __launch_bounds__(256) k_cell_rank(float4* __restrict__ cell_xyzf, unsigned int* __restrict__ cell_type,
                                                   unsigned int* __restrict__ cell_idx,
                                                   const unsigned int* __restrict__ scratch_idx,
                                                   const unsigned int* __restrict__ cell_start,
                                                   const unsigned int* __restrict__ particle_cell,
                                                   const float4* __restrict__ pos, const float* __restrict__ charge,
                                                   unsigned int ncells, int flag_charge, int flag_type,
                                                   const unsigned int* __restrict__ enable,
                                                   unsigned int* __restrict__ clean_counts = nullptr)
    {
    pdl_wait();
    pdl_trigger();
    if (enable && !*enable) return;
    unsigned int e = blockIdx.x * blockDim.x + threadIdx.x;
    // after k_cell_scan_scatter (which cannot clean the counts it shares between its CTAs): leave zeros for the next build
    if (clean_counts)
        for (unsigned int c = e; c < ncells; c += gridDim.x * blockDim.x) clean_counts[c] = 0u;
    unsigned int nbinned = cell_start[ncells];
    if (e >= nbinned) return;
    unsigned int n = scratch_idx[e];
    unsigned int cell = particle_cell[n];
    unsigned int s = cell_start[cell], t = cell_start[cell + 1];
    unsigned int rank = 0;
    for (unsigned int k = s; k < t; k++) rank += (scratch_idx[k] < n) ? 1u : 0u;
    float4 p = pos[n];
    float flag;
    if (flag_charge) flag = charge[n];
    else if (flag_type) flag = p.w;
    else flag = __int_as_float((int)n);
    cell_xyzf[s + rank] = make_float4(p.x, p.y, p.z, flag);
    if (cell_type) cell_type[s + rank] = (unsigned int)__float_as_int(p.w);
    if (cell_idx) cell_idx[s + rank] = n;
    }

// ---- small systems: the four phases in ONE cooperative launch ----------------------------------------------------------
// At 64k particles each of the four kernels above runs 3 us while the dependent-launch gap inside the conditional rebuild
// body is 5-8 us (profiles/r1n_*): the build is launch-bound.  One cooperative grid (<= one CTA of 512 threads per 512
// particles, all co-resident) does bin -> grid.sync -> scan (CTA 0) -> grid.sync -> scatter -> grid.sync -> rank with the
// same arithmetic and the same final order (the in-cell rank fixes it), so the result is bit-identical to the 4-kernel path.
constexpr unsigned int CF_BLOCK = 512;
constexpr unsigned int CF_MAX_CELLS = 4 * SCAN_TILE;

// Grid-wide barrier of a cooperative launch on two caller-owned words {arrivals, epoch} (b200_cell_list_args.d_fused_barrier):
// release atomic on arrival, the last CTA resets the count and publishes the next epoch, the others spin on it at GPU scope,
// one acquire fence per CTA (the same construction carries k_md_steps, md_steps_persist.cuh).
__device__ __forceinline__ void cf_grid_barrier(unsigned int* bar)
    {
    __syncthreads();
    if (threadIdx.x == 0)
        {
        unsigned int epoch, old;
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(epoch) : "l"(bar + 1) : "memory");
        asm volatile("atom.release.gpu.global.add.u32 %0, [%1], 1;" : "=r"(old) : "l"(bar) : "memory");
        if (old == gridDim.x - 1u)
            {
            asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(bar), "r"(0u) : "memory");
            asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(bar + 1), "r"(epoch + 1u) : "memory");
            }
        else
            {
            unsigned int now, spins = 0;
            do
                {
                asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(now) : "l"(bar + 1) : "memory");
                } while (now == epoch && ++spins < (1u << 22)); // bounded: a lost CTA must not hang the GPU
            }
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
        }
    __syncthreads();
    }

__global__ void __launch_bounds__(CF_BLOCK) k_cell_list_fused(
    unsigned int* __restrict__ cell_count, unsigned int* __restrict__ cell_start, unsigned int* __restrict__ particle_cell,
    unsigned int* __restrict__ particle_slot, unsigned int* __restrict__ scratch_idx, float4* __restrict__ cell_xyzf,
    unsigned int* __restrict__ cell_type, unsigned int* __restrict__ cell_idx, unsigned int* __restrict__ conditions,
    const float4* __restrict__ pos, const float* __restrict__ charge, unsigned int N, unsigned int n_ghost, Box box,
    uint3 dim, float3 ghost_width, int flag_charge, int flag_type, const unsigned int* __restrict__ enable,
    const unsigned int* __restrict__ d_N, unsigned int* barrier)
    {
    if (enable && !*enable) return; // uniform over the grid: nobody reaches a barrier
    const unsigned int ntot = N + n_ghost;
    const unsigned int ncells = dim.x * dim.y * dim.z;
    const unsigned int stride = gridDim.x * blockDim.x;
    const unsigned int tid = blockIdx.x * blockDim.x + threadIdx.x;
    // phase 1: bin (k_cell_bin)
    const unsigned int Nloc = d_N ? *d_N : N;
    for (unsigned int n = tid; n < ntot; n += stride)
        {
        if (n >= Nloc && n < N) // multi-GPU: unused local capacity
            {
            particle_cell[n] = B200_NOT_BINNED;
            continue;
            }
        const float4 p = pos[n];
        unsigned int cell = B200_NOT_BINNED, slot = 0;
        if (isnan(p.x) || isnan(p.y) || isnan(p.z))
            atomicMax(&conditions[1], n + 1);
        else
            {
            int3 b;
            if (particle_bin(make_float3(p.x, p.y, p.z), box, ghost_width, dim, b))
                {
                cell = ((unsigned int)b.z * dim.y + (unsigned int)b.y) * dim.x + (unsigned int)b.x;
                slot = atomicAdd(&cell_count[cell], 1u);
                }
            else if (n < N)
                atomicMax(&conditions[2], n + 1);
            }
        particle_cell[n] = cell;
        particle_slot[n] = slot;
        }
    cf_grid_barrier(barrier);
    // phase 2: exclusive scan of the occupancies by CTA 0 (k_cell_scan; 8 cells per thread and tile), self-cleaning counts
    if (blockIdx.x == 0)
        {
        __shared__ unsigned int warp_sums[CF_BLOCK / 32];
        __shared__ unsigned int carry_s;
        const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        if (threadIdx.x == 0) carry_s = 0;
        unsigned int max_occ = 0;
        __syncthreads();
        for (unsigned int base = 0; base < ncells; base += CF_BLOCK * 8)
            {
            const unsigned int i0 = base + threadIdx.x * 8;
            const unsigned int carry = carry_s;
            unsigned int c[8];
            unsigned int tsum = 0;
#pragma unroll
            for (int k = 0; k < 8; k++)
                {
                c[k] = (i0 + k < ncells) ? cell_count[i0 + k] : 0u;
                if (i0 + k < ncells) cell_count[i0 + k] = 0u;
                max_occ = max(max_occ, c[k]);
                tsum += c[k];
                }
            unsigned int incl = tsum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1)
                {
                const unsigned int v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
                }
            if (lane == 31) warp_sums[warp] = incl;
            __syncthreads();
            if (warp == 0)
                {
                const unsigned int w = lane < CF_BLOCK / 32 ? warp_sums[lane] : 0u;
                unsigned int wi = w;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1)
                    {
                    const unsigned int v = __shfl_up_sync(0xffffffffu, wi, o);
                    if (lane >= o) wi += v;
                    }
                if (lane < CF_BLOCK / 32) warp_sums[lane] = wi - w;
                if (lane == 31) carry_s = carry + wi;
                }
            __syncthreads();
            unsigned int excl = carry + warp_sums[warp] + incl - tsum;
#pragma unroll
            for (int k = 0; k < 8; k++)
                {
                if (i0 + k < ncells) cell_start[i0 + k] = excl;
                excl += c[k];
                }
            __syncthreads();
            }
        if (threadIdx.x == 0) cell_start[ncells] = carry_s;
        for (int o = 16; o > 0; o >>= 1) max_occ = max(max_occ, __shfl_xor_sync(0xffffffffu, max_occ, o));
        if (lane == 0 && max_occ) atomicMax(&conditions[0], max_occ);
        }
    cf_grid_barrier(barrier);
    // phase 3: scatter (k_cell_scatter).  cell_start was written by another CTA in this launch: no read-only path
    for (unsigned int n = tid; n < ntot; n += stride)
        {
        const unsigned int cell = particle_cell[n];
        if (cell == B200_NOT_BINNED) continue;
        scratch_idx[__ldcg(cell_start + cell) + particle_slot[n]] = n;
        }
    cf_grid_barrier(barrier);
    // phase 4: in-cell rank by particle index, final arrays (k_cell_rank)
    const unsigned int nbinned = __ldcg(cell_start + ncells);
    for (unsigned int e = tid; e < nbinned; e += stride)
        {
        const unsigned int n = __ldcg(scratch_idx + e);
        const unsigned int cell = __ldcg(particle_cell + n);
        const unsigned int s = __ldcg(cell_start + cell), t = __ldcg(cell_start + cell + 1);
        unsigned int rank = 0;
        for (unsigned int k = s; k < t; k++) rank += (__ldcg(scratch_idx + k) < n) ? 1u : 0u;
        const float4 p = pos[n];
        float flag;
        if (flag_charge) flag = charge[n];
        else if (flag_type) flag = p.w;
        else flag = __int_as_float((int)n);
        cell_xyzf[s + rank] = make_float4(p.x, p.y, p.z, flag);
        if (cell_type) cell_type[s + rank] = (unsigned int)__float_as_int(p.w);
        if (cell_idx) cell_idx[s + rank] = n;
        }
    }

// ---- small systems: the four phases in ONE launch of a single thread-block cluster (opt-in experiment, see below) --------
// The cooperative grid above loses because its three grid barriers and the cooperative launch cost what the launch gaps cost.
// A cluster's hardware barrier (barrier.cluster) is a few hundred ns, and 8-16 CTAs x 1024 threads are enough threads for
// <= 131k particles (the phases are chains of dependent loads / atomics, not bandwidth): each thread keeps the cell and slot
// of its <= CC_PER particles in registers from the bin phase to the scatter phase, every CTA scans ALL occupancies itself
// (ncells <= 8192: one 1024-thread block scan, result in shared memory), so only two cluster barriers remain -- after the
// counts and after the scatter.  Same arithmetic and the same final order as the 4-kernel path (the in-cell rank fixes it).
constexpr unsigned int CC_BLOCK = 1024;
constexpr unsigned int CC_PER = 8;       // particles per thread
constexpr unsigned int CC_CELLS_PER = 8; // cells per thread in the scan
constexpr unsigned int CC_MAX_CELLS = CC_BLOCK * CC_CELLS_PER;

__device__ __forceinline__ void cluster_barrier()
    {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    }

__global__ void __launch_bounds__(CC_BLOCK, 1) k_cell_list_cluster(
    unsigned int* __restrict__ cell_count, unsigned int* __restrict__ cell_start, unsigned int* __restrict__ particle_cell,
    unsigned int* __restrict__ particle_slot, unsigned int* scratch_idx, float4* __restrict__ cell_xyzf,
    unsigned int* __restrict__ cell_type, unsigned int* __restrict__ cell_idx, unsigned int* __restrict__ conditions,
    const float4* __restrict__ pos, const float* __restrict__ charge, unsigned int N, unsigned int n_ghost, Box box,
    uint3 dim, float3 ghost_width, int flag_charge, int flag_type, const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return; // uniform over the cluster: nobody reaches a barrier
    extern __shared__ unsigned int s_start[]; // [ncells + 1] exclusive scan of the occupancies
    __shared__ unsigned int warp_sums[33];
    const unsigned int ntot = N + n_ghost;
    const unsigned int ncells = dim.x * dim.y * dim.z;
    const unsigned int stride = gridDim.x * CC_BLOCK;
    const unsigned int tid = blockIdx.x * CC_BLOCK + threadIdx.x;
    // phase 1: bin (k_cell_bin); cell and slot stay in registers
    unsigned int cell_r[CC_PER], slot_r[CC_PER];
#pragma unroll
    for (unsigned int k = 0; k < CC_PER; k++)
        {
        const unsigned int n = tid + k * stride;
        cell_r[k] = B200_NOT_BINNED;
        slot_r[k] = 0;
        if (n < ntot)
            {
            const float4 p = pos[n];
            if (isnan(p.x) || isnan(p.y) || isnan(p.z))
                atomicMax(&conditions[1], n + 1);
            else
                {
                int3 b;
                if (particle_bin(make_float3(p.x, p.y, p.z), box, ghost_width, dim, b))
                    {
                    cell_r[k] = ((unsigned int)b.z * dim.y + (unsigned int)b.y) * dim.x + (unsigned int)b.x;
                    slot_r[k] = atomicAdd(&cell_count[cell_r[k]], 1u);
                    }
                else if (n < N)
                    atomicMax(&conditions[2], n + 1);
                }
            particle_cell[n] = cell_r[k];
            particle_slot[n] = slot_r[k];
            }
        }
    cluster_barrier();
    // phase 2: every CTA scans all occupancies (the counts live in L2: atomics); CTA 0 publishes cell_start
    const unsigned int i0 = threadIdx.x * CC_CELLS_PER;
    unsigned int c[CC_CELLS_PER];
    unsigned int tsum = 0, max_occ = 0;
#pragma unroll
    for (unsigned int k = 0; k < CC_CELLS_PER; k++)
        {
        c[k] = (i0 + k < ncells) ? __ldcg(cell_count + i0 + k) : 0u;
        tsum += c[k];
        max_occ = max(max_occ, c[k]);
        }
    unsigned int total;
    unsigned int excl = block_scan_1024(tsum, warp_sums, total);
#pragma unroll
    for (unsigned int k = 0; k < CC_CELLS_PER; k++)
        {
        if (i0 + k < ncells)
            {
            s_start[i0 + k] = excl;
            if (blockIdx.x == 0) cell_start[i0 + k] = excl;
            }
        excl += c[k];
        }
    if (threadIdx.x == 0)
        {
        s_start[ncells] = total;
        if (blockIdx.x == 0) cell_start[ncells] = total;
        }
    if (blockIdx.x == 0)
        {
        for (int o = 16; o > 0; o >>= 1) max_occ = max(max_occ, __shfl_xor_sync(0xffffffffu, max_occ, o));
        if ((threadIdx.x & 31) == 0 && max_occ) atomicMax(&conditions[0], max_occ);
        }
    __syncthreads();
    // phase 3: scatter (k_cell_scatter)
#pragma unroll
    for (unsigned int k = 0; k < CC_PER; k++)
        if (cell_r[k] != B200_NOT_BINNED) scratch_idx[s_start[cell_r[k]] + slot_r[k]] = tid + k * stride;
    cluster_barrier();
    // every CTA has read the counts: leave zeros behind for the next build (as k_cell_scan does)
    if (blockIdx.x == 0)
        {
#pragma unroll
        for (unsigned int k = 0; k < CC_CELLS_PER; k++)
            if (i0 + k < ncells) cell_count[i0 + k] = 0u;
        }
    // phase 4: in-cell rank by particle index, final arrays (k_cell_rank)
    const unsigned int nbinned = total;
#pragma unroll 2
    for (unsigned int k = 0; k < CC_PER; k++)
        {
        const unsigned int e = tid + k * stride;
        if (e >= nbinned) break;
        const unsigned int n = scratch_idx[e];
        const unsigned int cell = __ldcg(particle_cell + n);
        const unsigned int s = s_start[cell], t = s_start[cell + 1];
        unsigned int rank = 0;
        for (unsigned int q = s; q < t; q++) rank += (scratch_idx[q] < n) ? 1u : 0u;
        const float4 p = pos[n];
        float flag;
        if (flag_charge) flag = charge[n];
        else if (flag_type) flag = p.w;
        else flag = __int_as_float((int)n);
        cell_xyzf[s + rank] = make_float4(p.x, p.y, p.z, flag);
        if (cell_type) cell_type[s + rank] = (unsigned int)__float_as_int(p.w);
        if (cell_idx) cell_idx[s + rank] = n;
        }
    }

// ---- export to the reference's padded layout ------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_cell_export(unsigned int* __restrict__ cell_size, float4* __restrict__ xyzf,
                                                     float4* __restrict__ tdb, unsigned int* __restrict__ conditions,
                                                     const unsigned int* __restrict__ cell_start,
                                                     const float4* __restrict__ cell_xyzf,
                                                     const unsigned int* __restrict__ cell_idx,
                                                     const float4* __restrict__ pos, const float* __restrict__ diameter,
                                                     const unsigned int* __restrict__ body, unsigned int ncells,
                                                     unsigned int Nmax)
    {
    // one warp per cell: lanes stride the cell's entries
    unsigned int cell = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    unsigned int lane = threadIdx.x & 31;
    if (cell >= ncells) return;
    unsigned int s = cell_start[cell], size = cell_start[cell + 1] - s;
    if (lane == 0)
        {
        cell_size[cell] = size;
        if (size > Nmax) atomicMax(&conditions[0], size);
        }
    for (unsigned int o = lane; o < size && o < Nmax; o += 32)
        {
        xyzf[(size_t)cell * Nmax + o] = cell_xyzf[s + o];
        if (tdb)
            {
            unsigned int n = cell_idx[s + o];
            tdb[(size_t)cell * Nmax + o] = make_float4(pos[n].w, diameter ? diameter[n] : 1.0f,
                                                       __int_as_float(body ? (int)body[n] : -1), 0.0f);
            }
        }
    }

// ---- adjacency list on the device (CellList::initializeCellAdj) -----------------------------------------------------
__device__ __forceinline__ int sorted_unique_axis(int c, int r, int m, int* out)
    {
    // wrapped indices c-r..c+r mod m, sorted ascending and de-duplicated; at most 2r+1 <= 7 values
    int n = 0;
    for (int d = -r; d <= r; d++)
        {
        int w = (c + d) % m;
        if (w < 0) w += m;
        int k = 0;
        while (k < n && out[k] < w) k++;
        if (k < n && out[k] == w) continue;
        for (int q = n; q > k; q--) out[q] = out[q - 1];
        out[k] = w;
        n++;
        }
    return n;
    }

__global__ void __launch_bounds__(128) k_cell_adj(unsigned int* __restrict__ cell_adj, uint3 dim, int radius,
                                                  unsigned int n_adj)
    {
    unsigned int cell = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int ncells = dim.x * dim.y * dim.z;
    if (cell >= ncells) return;
    int i = cell % dim.x, j = (cell / dim.x) % dim.y, k = cell / (dim.x * dim.y);
    int xi[7], yj[7], zk[7];
    int nx = sorted_unique_axis(i, radius, dim.x, xi);
    int ny = sorted_unique_axis(j, radius, dim.y, yj);
    int nz = sorted_unique_axis(k, radius, dim.z, zk);
    // id = (k*dimy + j)*dimx + i is lexicographic in (k,j,i): the sorted unique id list is the product of the sorted axes
    unsigned int a = 0;
    for (int c = 0; c < nz; c++)
        for (int b = 0; b < ny; b++)
            for (int q = 0; q < nx; q++)
                if (a < n_adj)
                    cell_adj[(size_t)cell * n_adj + a++] = ((unsigned int)zk[c] * dim.y + yj[b]) * dim.x + xi[q];
    }
} // namespace b200

using namespace b200;

extern "C" {

void b200_box_init(b200_box* box, float Lx, float Ly, float Lz)
    {
    const float L[3] = {Lx, Ly, Lz};
    for (int d = 0; d < 3; d++)
        {
        box->L[d] = L[d];
        box->hi[d] = L[d] / 2.0f;
        box->lo[d] = -box->hi[d];
        box->Linv[d] = 1.0f / L[d];
        box->periodic[d] = 1;
        }
    box->xy = box->xz = box->yz = 0.0f;
    box->_pad = 0;
    }

void b200_box_init_lohi(b200_box* box, const float lo[3], const float hi[3], const unsigned char periodic[3])
    {
    for (int d = 0; d < 3; d++)
        {
        box->lo[d] = lo[d];
        box->hi[d] = hi[d];
        box->Linv[d] = 1.0f / (hi[d] - lo[d]);
        box->L[d] = hi[d] - lo[d];
        box->periodic[d] = periodic[d];
        }
    box->xy = box->xz = box->yz = 0.0f;
    box->_pad = 0;
    }

void b200_cell_list_dimensions(const b200_box* box, float nominal_width, unsigned int multiple,
                               float ghost_layer_width, b200_dim3* dim, float ghost_width[3])
    {
    // hoomd/CellList.cc:69-131 (3D).  L = nearest plane distance = box length for an orthorhombic box.
    unsigned int d[3];
    if (multiple == 0) multiple = 1;
    for (int a = 0; a < 3; a++)
        {
        unsigned int v = (unsigned int)(box->L[a] / nominal_width);
        d[a] = (v / multiple) * multiple;
        ghost_width[a] = 0.0f;
        }
    if (ghost_layer_width > 0.0f)
        {
        for (int a = 0; a < 3; a++)
            if (!box->periodic[a]) ghost_width[a] = ghost_layer_width;
        for (int a = 0; a < 3; a++)
            if (!box->periodic[a])
                {
                float cell_size = box->L[a] / (float)d[a];
                d[a] += (unsigned int)(int)ceilf(ghost_width[a] / cell_size);
                }
        }
    for (int a = 0; a < 3; a++)
        if (d[a] == 0) d[a] = 1;
    dim->x = d[0]; dim->y = d[1]; dim->z = d[2];
    }

} // extern "C" (reopened below)

namespace b200
{
bool pdl_enabled()
    {
    static const char* env = getenv("HOOMD_B200_PDL");
    static const bool on = env ? atoi(env) != 0 : false;
    return on;
    }

// co-resident CTAs of the cooperative build on the current device (0: cooperative launch unavailable or switched off)
static unsigned int cell_fused_max_ctas()
    {
    static int cached = -1;
    if (cached < 0)
        {
        cached = 0;
        // Measured SLOWER than the 4-kernel build inside the conditional graph body, with cooperative_groups' grid.sync() (LJ
        // 64k: 31.5k vs 32.9k timesteps/s, profiles/r1p_*) and with the microsecond barrier above as well (34.1k vs 35.4k, cell
        // list 27.6 vs 24.9 us, profiles/r2q_*): the four phases are chains of dependent atomics / loads of 64k threads, not
        // launch gaps, and one CTA per 512 particles leaves SMs idle.  Kept behind HOOMD_B200_CELL_FUSED=1 for A/B runs.
        const char* env = getenv("HOOMD_B200_CELL_FUSED");
        int dev = 0, coop = 0, sms = 0, per_sm = 0;
        if (env && atoi(env) != 0 && cudaGetDevice(&dev) == cudaSuccess
            && cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev) == cudaSuccess && coop
            && cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess
            && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_cell_list_fused, (int)CF_BLOCK, 0) == cudaSuccess)
            cached = sms * per_sm;
        }
    return (unsigned int)cached;
    }

static bool cell_use_fused(unsigned int ncells, unsigned int ntot, bool have_barrier)
    {
    // launch-bound sizes only: beyond a few hundred thousand particles the kernels dwarf the launch gaps
    return have_barrier && ntot > 0 && ncells <= CF_MAX_CELLS && ntot <= 262144u && cell_fused_max_ctas() > 0;
    }

// CTAs of the one-cluster build on the current device (0: not requested or clusters unavailable): 16 where the device
// schedules a non-portable cluster of that size, else 8
static unsigned int cell_cluster_size()
    {
    static int cached = -1;
    if (cached < 0)
        {
        cached = 0;
        // Measured SLOWER than the 4-kernel build as well (LJ 64k, cell list alone: 32.0 us with 16 CTAs, 44.3 us with 8,
        // against 25.7 us; 32.9k / 31.3k vs 35.0k timesteps/s, profiles/r2x_*): 16 SMs do not issue 64k returning atomics
        // and 1M rank loads faster than 148 SMs pay four launch gaps.  Opt-in: HOOMD_B200_CELL_CLUSTER=8 | 16.
        const char* env = getenv("HOOMD_B200_CELL_CLUSTER");
        if (!env || atoi(env) < 8) return 0u;
        const int want = atoi(env) >= 16 ? 16 : 8;
        for (int cs = want; cs >= 8 && !cached; cs /= 2)
            {
            if (cs > 8
                && cudaFuncSetAttribute(k_cell_list_cluster, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess)
                {
                cudaGetLastError();
                continue;
                }
            cudaLaunchConfig_t cfg = {};
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = (unsigned int)cs;
            attr[0].val.clusterDim.y = 1;
            attr[0].val.clusterDim.z = 1;
            cfg.gridDim = dim3((unsigned int)cs);
            cfg.blockDim = dim3(CC_BLOCK);
            cfg.dynamicSmemBytes = sizeof(unsigned int) * (CC_MAX_CELLS + 1);
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            int nclusters = 0;
            if (cudaOccupancyMaxActiveClusters(&nclusters, k_cell_list_cluster, &cfg) == cudaSuccess && nclusters > 0)
                cached = cs;
            else
                cudaGetLastError();
            }
        }
    return (unsigned int)cached;
    }

// single: one domain (no live particle count on the device, no ghost shifts)
static bool cell_use_cluster(unsigned int ncells, unsigned int ntot, bool single)
    {
    if (!single || ntot == 0 || ncells > CC_MAX_CELLS) return false;
    const unsigned int cs = cell_cluster_size();
    return cs > 0 && ntot <= cs * CC_BLOCK * CC_PER;
    }

// scan folded into the scatter launch: small grids with enough particles to make the extra launch the larger cost
static bool cell_use_scan_scatter(unsigned int ncells, unsigned int ntot)
    {
    static const char* env = getenv("HOOMD_B200_CELL_SCANSCATTER");
    static const bool on = env ? atoi(env) != 0 : true;
    return on && ntot > 0 && ncells <= SS_MAX_CELLS && ntot <= 262144u;
    }

// launches per cell-list build: the multi-CTA scan (one launch more) pays off beyond a few tiles
unsigned int cell_list_kernels(unsigned int ncells, unsigned int ntot, bool have_barrier)
    {
    if (cell_use_cluster(ncells, ntot, have_barrier)) return 1u;
    if (cell_use_fused(ncells, ntot, have_barrier)) return 1u;
    if (cell_use_scan_scatter(ncells, ntot)) return 3u;
    const unsigned int ntiles = (ncells + SCAN_TILE - 1) / SCAN_TILE;
    return (ntiles > 4 && ntiles <= ntot) ? 5u : 4u;
    }

// enable == NULL: unconditional (driver entry point, zeroes cell_count itself).  enable != NULL: every kernel returns
// at once unless *enable != 0, and cell_count must already be zero (k_cell_scan leaves it zeroed).
cudaError_t launch_cell_list(const b200_cell_list_args* a, const unsigned int* enable, cudaStream_t stream,
                             const unsigned int* d_N)
    {
    if (!box_is_orthorhombic(a->box)) return cudaErrorNotSupported;
    const unsigned int ntot = a->N + a->n_ghost;
    const unsigned int ncells = a->dim.x * a->dim.y * a->dim.z;
    Box box = make_box(a->box);
    uint3 dim = make_uint3(a->dim.x, a->dim.y, a->dim.z);
    float3 gw = make_float3(a->ghost_width[0], a->ghost_width[1], a->ghost_width[2]);
    if (!enable) cudaMemsetAsync(a->d_cell_count, 0, sizeof(unsigned int) * ncells, stream);
    if (cell_use_cluster(ncells, ntot, !d_N && !a->d_ghost_dir))
        {
        const unsigned int cs = cell_cluster_size();
        cudaLaunchConfig_t cfg = {};
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = cs;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.gridDim = dim3(cs);
        cfg.blockDim = dim3(CC_BLOCK);
        cfg.dynamicSmemBytes = sizeof(unsigned int) * (ncells + 1);
        cfg.stream = stream;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        return cudaLaunchKernelEx(&cfg, k_cell_list_cluster, a->d_cell_count, a->d_cell_start, a->d_particle_cell,
                                  a->d_particle_slot, a->d_scratch_idx, (float4*)a->d_cell_xyzf, a->d_cell_type,
                                  a->d_cell_idx, a->d_conditions, (const float4*)a->d_pos, (const float*)a->d_charge, a->N,
                                  a->n_ghost, box, dim, gw, (int)a->flag_charge, (int)a->flag_type, enable);
        }
    const bool fused = cell_use_fused(ncells, ntot, a->d_fused_barrier != nullptr) && !a->d_ghost_dir; // (bins ghosts where they are)
    if (fused)
        {
        unsigned int grid = (ntot + CF_BLOCK - 1) / CF_BLOCK;
        if (grid > cell_fused_max_ctas()) grid = cell_fused_max_ctas();
        unsigned int N = a->N, n_ghost = a->n_ghost;
        const float4* pos = (const float4*)a->d_pos;
        float4* xyzf = (float4*)a->d_cell_xyzf;
        int flag_charge = a->flag_charge, flag_type = a->flag_type;
        void* args[] = {(void*)&a->d_cell_count, (void*)&a->d_cell_start, (void*)&a->d_particle_cell,
                        (void*)&a->d_particle_slot, (void*)&a->d_scratch_idx, &xyzf, (void*)&a->d_cell_type,
                        (void*)&a->d_cell_idx, (void*)&a->d_conditions, &pos, (void*)&a->d_charge, &N, &n_ghost, &box,
                        &dim, &gw, &flag_charge, &flag_type, &enable, &d_N, (void*)&a->d_fused_barrier};
        return cudaLaunchCooperativeKernel((const void*)k_cell_list_fused, dim3(grid), dim3(CF_BLOCK), args, 0, stream);
        }
    if (ntot)
        k_cell_bin<<<(ntot + 255) / 256, 256, 0, stream>>>(a->d_cell_count, a->d_particle_cell, a->d_particle_slot,
                                                            a->d_conditions, (const float4*)a->d_pos, a->N, a->n_ghost,
                                                            box, dim, gw, enable, d_N, a->d_ghost_dir, a->d_ghost_shift);
    const unsigned int ntiles = (ncells + SCAN_TILE - 1) / SCAN_TILE;
    // device-driven rebuild: the dependent kernels are scheduled behind their predecessor (see launch_chain)
    const bool pdl = enable && pdl_enabled();
    const bool scan_scatter = cell_use_scan_scatter(ncells, ntot);
    if (scan_scatter)
        launch_chain(k_cell_scan_scatter, dim3((ntot + SS_BLOCK - 1) / SS_BLOCK), dim3(SS_BLOCK), 0, stream, pdl,
                     a->d_cell_start, (const unsigned int*)a->d_cell_count, ncells, a->d_conditions, a->d_scratch_idx,
                     (const unsigned int*)a->d_particle_cell, (const unsigned int*)a->d_particle_slot, ntot, enable);
    else if (cell_list_kernels(ncells, ntot, false) == 5)
        {
        // tile totals live in d_scratch_idx, which is not needed before k_cell_scatter
        launch_chain(k_cell_scan_tiles, dim3(ntiles), dim3(1024), 0, stream, pdl, a->d_cell_start, a->d_cell_count, ncells,
                     a->d_scratch_idx, a->d_conditions, enable);
        launch_chain(k_cell_scan_fix, dim3(ntiles), dim3(1024), 0, stream, pdl, a->d_cell_start, ncells, a->d_scratch_idx,
                     enable);
        }
    else
        launch_chain(k_cell_scan, dim3(1), dim3(1024), 0, stream, pdl, a->d_cell_start, a->d_cell_count, ncells,
                     a->d_conditions, enable);
    if (ntot)
        {
        if (!scan_scatter)
            launch_chain(k_cell_scatter, dim3((ntot + 255) / 256), dim3(256), 0, stream, pdl, a->d_scratch_idx,
                         a->d_cell_start, a->d_particle_cell, a->d_particle_slot, ntot, enable);
        launch_chain(k_cell_rank, dim3((ntot + 255) / 256), dim3(256), 0, stream, pdl, (float4*)a->d_cell_xyzf,
                     a->d_cell_type, a->d_cell_idx, a->d_scratch_idx, a->d_cell_start, a->d_particle_cell,
                     (const float4*)a->d_pos, a->d_charge, ncells, a->flag_charge, a->flag_type, enable,
                     scan_scatter ? a->d_cell_count : (unsigned int*)nullptr);
        }
    return cudaGetLastError();
    }
} // namespace b200

extern "C" {

int b200_compute_cell_list(const b200_cell_list_args* a, b200_stream_t stream)
    {
    return (int)b200::launch_cell_list(a, nullptr, (cudaStream_t)stream, nullptr);
    }

int b200_cell_list_export_padded(unsigned int* d_cell_size, float* d_xyzf, float* d_tdb, unsigned int* d_conditions,
                                 const unsigned int* d_cell_start, const float* d_cell_xyzf,
                                 const unsigned int* d_cell_idx, const float* d_pos, const float* d_diameter,
                                 const unsigned int* d_body, unsigned int ncells, unsigned int Nmax,
                                 b200_stream_t stream)
    {
    if (d_tdb && !d_cell_idx) return (int)cudaErrorInvalidValue;
    unsigned int threads = ncells * 32;
    k_cell_export<<<(threads + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        d_cell_size, (float4*)d_xyzf, (float4*)d_tdb, d_conditions, d_cell_start, (const float4*)d_cell_xyzf,
        d_cell_idx, (const float4*)d_pos, d_diameter, d_body, ncells, Nmax);
    return (int)cudaGetLastError();
    }

int b200_cell_list_build_adjacency(unsigned int* d_cell_adj, b200_dim3 dim, unsigned int radius, unsigned int n_adj,
                                   b200_stream_t stream)
    {
    if (radius > 3) return (int)cudaErrorInvalidValue;
    unsigned int ncells = dim.x * dim.y * dim.z;
    k_cell_adj<<<(ncells + 127) / 128, 128, 0, (cudaStream_t)stream>>>(d_cell_adj, make_uint3(dim.x, dim.y, dim.z),
                                                                        (int)radius, n_adj);
    return (int)cudaGetLastError();
    }

} // extern "C"
