// Multi-GPU communication kernels (see comm.cuh).  Everything here runs on the rank's own stream; peers are reached
// through pointers into their IPC-mapped slabs (plain st.global over NVLink/NVSwitch).
#include "comm.cuh"
#include "decide.cuh"
#include <algorithm>

namespace b200
{
__device__ __forceinline__ bool warp_barrier(const MgConst& c, MgState* ms, int kind, bool payload)
    {
    return mg_warp_barrier(c.world, c.rank, c.flags, ms, kind, payload);
    }

__global__ void __launch_bounds__(32) k_mg_step_barrier(const __grid_constant__ MgConst c, MgState* ms, StepState* st,
                                                        cudaGraphConditionalHandle handle, int use_cond,
                                                        unsigned long long* nvt_acc)
    {
    bool moved;
    if (nvt_acc)
        {
        // TwoStepNVTMTK under decomposition: the thermostat needs sum(m v^2) of ALL ranks.  The step kernel left this rank's
        // share in the fixed-point accumulators of the parity the NEXT launch reads (its ts_step_two = the current
        // timestep); the barrier adds up the ranks' totals and leaves the world total in slot 0, so md_step_nvt_begin of
        // every rank advances its copy of the thermostat with the same bits.
        unsigned long long* acc = nvt_acc + (size_t)(st->timestep & 1u) * NVT_SLOTS * NVT_SLOT_STRIDE;
        unsigned long long v = threadIdx.x < NVT_SLOTS ? acc[threadIdx.x * NVT_SLOT_STRIDE] : 0ull;
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        unsigned long long total = 0ull;
        moved = mg_warp_barrier(c.world, c.rank, c.flags, ms, 0, st->moved != 0u, v, &total);
        if (threadIdx.x < NVT_SLOTS) acc[threadIdx.x * NVT_SLOT_STRIDE] = threadIdx.x == 0 ? total : 0ull;
        __syncwarp();
        }
    else
        // the vote: did any particle of any rank move more than r_buff/2 (hoomd/md/NeighborListGPU.cc:117-122)
        moved = warp_barrier(c, ms, 0, st->moved != 0u);
    if (threadIdx.x == 0)
        {
        st->moved = moved ? 1u : 0u;
        if (st->steps_left && !st->abort)
            decide_body(st);
        else
            {
            st->need_rebuild = 0u;
            st->moved = 0u;
            }
#if (CUDART_VERSION >= 12040)
        if (use_cond) cudaGraphSetConditional(handle, st->need_rebuild);
#endif
        }
    }

__global__ void __launch_bounds__(32) k_mg_barrier(const __grid_constant__ MgConst c, MgState* ms, int kind,
                                                   const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    warp_barrier(c, ms, kind, false);
    }

// last kernel of the rebuild bracket: did the neighbor list of ANY rank overflow Nmax?  Then no rank may step on (the
// host grows Nmax and the chunk resumes from this very step): the reference's overflow loop
// (hoomd/md/NeighborList.cc:226-244) made collective.
__global__ void __launch_bounds__(32) k_mg_rebuild_vote(const __grid_constant__ MgConst c, MgState* ms, StepState* st,
                                                        int buffer, const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    const bool any = warp_barrier(c, ms, 3, st->overflow != 0u);
    if (threadIdx.x == 0 && any) st->abort = 1u + (unsigned int)buffer;
    }

// ---- rebuild bracket, part 1: migration ---------------------------------------------------------------------------------
// reset the per-rebuild counters, blank this rank's ghost region of the current position buffer (slots no neighbor
// fills this time must not be binned: a far-away position is ignored by the cell list like any out-of-range ghost,
// hoomd/CellList.cc:494-498)
__global__ void __launch_bounds__(256) k_mg_begin(const __grid_constant__ MgConst c, MgState* ms, float4* __restrict__ pos,
                                                  unsigned int* __restrict__ chunk_flag, unsigned int nchunks,
                                                  const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    const unsigned int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < c.Gcap) pos[c.Ncap + t] = make_float4(1.0e30f, 1.0e30f, 1.0e30f, 0.0f);
    if (t < nchunks) chunk_flag[t] = 0u;
    if (t < MG_NDIR)
        {
        ms->send_cnt[t] = 0u;
        ms->gsend_cnt[t] = 0u;
        }
    if (t == 0)
        {
        ms->n_leave = 0u;
        ms->n_low_holes = 0u;
        ms->n_moved_down = 0u;
        ms->rebuild_id += 1u;
        }
    }

// which neighbor owns the particle now (DomainDecomposition::placeParticle restricted to the 26 adjacent domains);
// migrants are written straight into the new owner's mailbox, already translated into its frame (the periodic shift is
// the `x - L` / `x + L` of BoxDim::wrap, hoomd/BoxDim.h:366-432, with the image flag adjusted accordingly)
__global__ void __launch_bounds__(256) k_mg_migrate_mark(const __grid_constant__ MgConst c, const MgLocal l,
                                                         const float4* __restrict__ pos,
                                                         const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int N = l.st->N;
    if (i >= N) return;
    const float4 p = pos[i];
    const float x[3] = {p.x, p.y, p.z};
    int d[3] = {0, 0, 0};
#pragma unroll
    for (int a = 0; a < 3; a++)
        if (c.split[a])
            {
            if (x[a] < c.lo[a]) d[a] = -1;
            else if (x[a] >= c.hi[a]) d[a] = 1;
            }
    const int code = (d[2] + 1) * 9 + (d[1] + 1) * 3 + (d[0] + 1);
    if (code == MG_DIR_STAY) return;
    const unsigned int rid = l.ms->rebuild_id;
    const unsigned int slot = atomicAdd(&l.ms->send_cnt[code], 1u);
    if (slot < c.mig_cap && c.dir_rank[code] >= 0)
        {
        MigElem e;
        e.pos = make_float4(__fadd_rn(p.x, c.shift[code][0]), __fadd_rn(p.y, c.shift[code][1]),
                            __fadd_rn(p.z, c.shift[code][2]), p.w);
        e.vel = l.vel[i];
        e.ax = l.accel[3 * i + 0];
        e.ay = l.accel[3 * i + 1];
        e.az = l.accel[3 * i + 2];
        e.ix = l.image[3 * i + 0] + (c.shift[code][0] < 0.0f ? 1 : (c.shift[code][0] > 0.0f ? -1 : 0));
        e.iy = l.image[3 * i + 1] + (c.shift[code][1] < 0.0f ? 1 : (c.shift[code][1] > 0.0f ? -1 : 0));
        e.iz = l.image[3 * i + 2] + (c.shift[code][2] < 0.0f ? 1 : (c.shift[code][2] > 0.0f ? -1 : 0));
        e.tag = l.tag[i];
        e._pad = 0u;
        c.peer_mig[code][slot] = e;
        }
    else
        atomicOr(&l.ms->error, 1u);
    const unsigned int h = atomicAdd(&l.ms->n_leave, 1u);
    l.holes[h] = i;
    l.hole_mark[i] = rid;
    }

// publish the counts to the receivers, then barrier 1 (all mailboxes complete)
__global__ void __launch_bounds__(32) k_mg_migrate_publish(const __grid_constant__ MgConst c, MgState* ms,
                                                           const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    const unsigned int d = threadIdx.x;
    if (d < MG_NDIR && d != MG_DIR_STAY && c.dir_rank[d] >= 0)
        {
        const unsigned int n = min(ms->send_cnt[d], c.mig_cap);
        *c.peer_mig_cnt[d] = n;
        atomicMax(&ms->max_send, n);
        }
    warp_barrier(c, ms, 1, false);
    // prefix of what arrived; new local count
    unsigned int cnt = 0;
    if (d < MG_NDIR && d != MG_DIR_STAY && c.dir_rank[d] >= 0) cnt = *((volatile unsigned int*)&c.my_mig_cnt[d]);
    unsigned int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
        {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((int)d >= o) incl += t;
        }
    if (d < MG_NDIR) ms->recv_off[d] = incl - cnt;
    if (d == MG_NDIR - 1) ms->recv_off[MG_NDIR] = incl;
    }

__global__ void __launch_bounds__(32) k_mg_migrate_count(const __grid_constant__ MgConst c, const MgLocal l,
                                                         const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    if (threadIdx.x != 0) return;
    const unsigned int N = l.st->N, L = l.ms->n_leave, R = l.ms->recv_off[MG_NDIR];
    unsigned int N_new = N - L + R;
    if (N_new > c.Ncap)
        {
        atomicOr(&l.ms->error, 4u);
        N_new = c.Ncap;
        }
    l.ms->N_new = N_new;
    atomicMax(&l.ms->max_N, N_new);
    }

// received particle g goes to the g-th hole while holes last, then to the end of the arrays
__global__ void __launch_bounds__(128) k_mg_migrate_unpack(const __grid_constant__ MgConst c, const MgLocal l,
                                                           float4* __restrict__ pos,
                                                           const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    const unsigned int d = blockIdx.y;
    const unsigned int r = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int cnt = l.ms->recv_off[d + 1] - l.ms->recv_off[d];
    if (r >= cnt) return;
    const unsigned int N = l.st->N, L = l.ms->n_leave;
    const unsigned int g = l.ms->recv_off[d] + r;
    unsigned int slot;
    if (g < L)
        {
        slot = l.holes[g];
        l.hole_mark[slot] = 0u;
        }
    else
        slot = N + (g - L);
    if (slot >= c.Ncap) return; // capacity error already flagged
    const MigElem e = c.my_mig[(size_t)d * c.mig_cap + r];
    pos[slot] = e.pos;
    l.vel[slot] = e.vel;
    l.accel[3 * slot + 0] = e.ax;
    l.accel[3 * slot + 1] = e.ay;
    l.accel[3 * slot + 2] = e.az;
    l.image[3 * slot + 0] = e.ix;
    l.image[3 * slot + 1] = e.iy;
    l.image[3 * slot + 2] = e.iz;
    l.tag[slot] = e.tag;
    }

// more left than arrived: holes below the new count are listed ...
__global__ void __launch_bounds__(256) k_mg_low_holes(const MgLocal l, const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    const unsigned int L = l.ms->n_leave, R = l.ms->recv_off[MG_NDIR], N_new = l.ms->N_new;
    for (unsigned int h = R + blockIdx.x * blockDim.x + threadIdx.x; h < L; h += gridDim.x * blockDim.x)
        {
        const unsigned int idx = l.holes[h];
        if (idx < N_new) l.low_holes[atomicAdd(&l.ms->n_low_holes, 1u)] = idx;
        }
    }

// ... and filled with the particles that sit at or above it
__global__ void __launch_bounds__(256) k_mg_move_high(const MgLocal l, float4* __restrict__ pos,
                                                      const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    const unsigned int N = l.st->N, N_new = l.ms->N_new, rid = l.ms->rebuild_id;
    for (unsigned int idx = N_new + blockIdx.x * blockDim.x + threadIdx.x; idx < N; idx += gridDim.x * blockDim.x)
        {
        if (l.hole_mark[idx] == rid) continue; // vacated
        const unsigned int dst = l.low_holes[atomicAdd(&l.ms->n_moved_down, 1u)];
        pos[dst] = pos[idx];
        l.vel[dst] = l.vel[idx];
        for (int k = 0; k < 3; k++)
            {
            l.accel[3 * dst + k] = l.accel[3 * idx + k];
            l.image[3 * dst + k] = l.image[3 * idx + k];
            }
        l.tag[dst] = l.tag[idx];
        }
    }

__global__ void k_mg_commit(const MgLocal l, const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    l.st->N = l.ms->N_new;
    }

// ---- rebuild bracket, part 2: ghost plan ------------------------------------------------------------------------------------
// A particle within r_ghost of a face of a split axis is a ghost of the neighbor across that face; faces combine into
// edges and corners (<= 7 destinations).  Slot s of direction d on the receiver is fed by local particle plan_src[...]
// until the next rebuild (CommunicatorGPU::exchangeGhosts builds the same plan with a pack/sort/scan pipeline).
__global__ void __launch_bounds__(256) k_mg_ghost_plan(const __grid_constant__ MgConst c, const MgLocal l,
                                                       const float4* __restrict__ pos,
                                                       const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int lane = threadIdx.x & 31;
    const bool active = i < l.st->N;
    const float4 p = active ? pos[i] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    const float x[3] = {p.x, p.y, p.z};
    bool nlo[3], nhi[3];
    bool any = false, near = false;
    // a particle can have a ghost in its neighbor list only when it sits within r_list = r_ghost of a split face at build
    // time; slightly widened (FP32 rounding of the shifted ghost positions), it marks its chunk as a boundary chunk
    const float wide = c.r_ghost * 1.00001f + 1.0e-4f;
#pragma unroll
    for (int a = 0; a < 3; a++)
        {
        nlo[a] = active && c.split[a] && (x[a] < c.lo[a] + c.r_ghost);
        nhi[a] = active && c.split[a] && (x[a] >= c.hi[a] - c.r_ghost);
        any |= nlo[a] | nhi[a];
        near |= active && c.split[a] && (x[a] <= c.lo[a] + wide || x[a] >= c.hi[a] - wide);
        }
    if (near) l.chunk_flag[i / l.ppb] = 1u; // same-value stores
    if (!__any_sync(0xffffffffu, any))
        {
        if (active) l.pplan_n[i] = 0; // interior warp
        return;
        }
    const unsigned int ptag = active ? l.tag[i] : 0u;
    unsigned int nplan = 0;
    // Slots are handed out per warp (one atomic per warp and direction, lanes ranked by ballot): the particles of a
    // warp are neighbors along the space-filling curve, so their ghost copies land next to each other on the receiver
    // and its position gathers keep their locality.
    for (int d = 0; d < MG_NDIR; d++)
        {
        if (c.seg_cap[d] == 0u) continue; // unused direction (uniform)
        const int dx = d % 3 - 1, dy = (d / 3) % 3 - 1, dz = d / 9 - 1;
        const bool send = !((dx < 0 && !nlo[0]) || (dx > 0 && !nhi[0]) || (dy < 0 && !nlo[1]) || (dy > 0 && !nhi[1])
                            || (dz < 0 && !nlo[2]) || (dz > 0 && !nhi[2]));
        const unsigned int mask = __ballot_sync(0xffffffffu, send);
        if (!mask) continue;
        unsigned int base = 0;
        if (lane == (unsigned int)(__ffs(mask) - 1)) base = atomicAdd(&l.ms->gsend_cnt[d], (unsigned int)__popc(mask));
        base = __shfl_sync(0xffffffffu, base, __ffs(mask) - 1);
        if (send)
            {
            const unsigned int slot = base + __popc(mask & ((1u << lane) - 1u));
            if (slot < c.seg_cap[d])
                {
                l.plan_src[c.seg_off[d] + slot] = i;
                c.peer_tag[d][c.Ncap + c.seg_off[d] + slot] = ptag;
                l.pplan[(size_t)nplan * c.Ncap + i] = ((unsigned int)d << 27) | (c.seg_off[d] + slot); // <= 7 destinations
                nplan++;
                }
            else
                atomicOr(&l.ms->error, 2u);
            }
        }
    if (active) l.pplan_n[i] = (unsigned char)nplan;
    }

// boundary chunks first: ordered compaction of the chunk flags by one CTA (a few thousand chunks: per-thread ranges + a block
// scan), so consecutive CTAs of k_md_step still work on consecutive chunks
__global__ void __launch_bounds__(1024) k_mg_cta_order(const MgLocal l, const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    __shared__ unsigned int s_warp[32];
    __shared__ unsigned int s_total;
    const unsigned int n = l.nchunks, per = (n + 1023u) / 1024u;
    const unsigned int b = threadIdx.x * per, e = min(n, b + per);
    unsigned int cnt = 0;
    for (unsigned int k = b; k < e; k++) cnt += l.chunk_flag[k] ? 1u : 0u;
    // exclusive block scan of cnt
    unsigned int incl = cnt;
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
        {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((int)lane >= o) incl += t;
        }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0)
        {
        unsigned int w = s_warp[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
            {
            const unsigned int t = __shfl_up_sync(0xffffffffu, wi, o);
            if ((int)lane >= o) wi += t;
            }
        s_warp[lane] = wi - w;
        if (lane == 31) s_total = wi;
        }
    __syncthreads();
    const unsigned int nb = s_total;
    unsigned int ob = s_warp[warp] + incl - cnt; // boundary chunks before this thread's range
    unsigned int oi = nb + (b < n ? b : n) - ob; // interior chunks before it, placed behind all boundary chunks
    for (unsigned int k = b; k < e; k++)
        {
        if (l.chunk_flag[k]) l.cta_order[ob++] = k;
        else l.cta_order[oi++] = k;
        }
    if (threadIdx.x == 0) l.ms->n_bchunks = nb;
    }

// every MD step: one store per (boundary particle, destination) into the neighbor's position buffer
__global__ void __launch_bounds__(256) k_mg_ghost_send(const __grid_constant__ MgConst c, const MgLocal l,
                                                       const float4* __restrict__ pos,
                                                       const unsigned char* __restrict__ slot_dir, int buffer,
                                                       const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    const unsigned int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= c.Gcap) return;
    const unsigned int d = slot_dir[t];
    const unsigned int slot = t - c.seg_off[d];
    const unsigned int n = min(l.ms->gsend_cnt[d], c.seg_cap[d]);
    if (slot >= n) return;
    const unsigned int i = l.plan_src[t];
    // raw coordinates: the receiver bins the ghost at its periodic image (b200_cell_list_args.d_ghost_shift) and takes
    // minimum images in the global box, so a ghost pair is evaluated with the single-domain arithmetic
    float4* dst = buffer ? c.peer_posB[d] : c.peer_posA[d];
    dst[c.Ncap + t] = pos[i];
    if (slot == 0) atomicMax(&l.ms->max_gsend, n);
    }

// ---- tag -> index map over locals AND ghosts (bonded forces / exclusions under decomposition) ---------------------------------
// Bond and exclusion tables stay keyed by GLOBAL tag on every rank (replicated, like the reference's BondData before
// GroupCommunicator scatters it); a bond partner that lives on another rank is reached through its ghost copy.  Unused local
// slots [N, Ncap) get the tag NO_TAG so that the tag-keyed kernels skip them.
__global__ void __launch_bounds__(256) k_mg_rtag_clear(unsigned int* __restrict__ rtag, unsigned int ntags,
                                                       const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    const unsigned int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < ntags) rtag[t] = 0xffffffffu;
    }

__global__ void __launch_bounds__(256) k_mg_rtag_fill(const __grid_constant__ MgConst c, const MgLocal l,
                                                      const float4* __restrict__ pos, unsigned int* __restrict__ rtag,
                                                      unsigned int ntags, const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= c.Ncap + c.Gcap) return;
    const unsigned int N = l.st->N;
    if (i < c.Ncap)
        {
        if (i >= N)
            {
            l.tag[i] = 0xffffffffu;
            return;
            }
        }
    else if (!(pos[i].x < 1.0e29f))
        return; // blank ghost slot
    const unsigned int t = l.tag[i];
    // locals win over ghost copies of the same particle (none exist for legal domain widths); among copies any one serves:
    // they hold the same raw coordinates
    if (t < ntags && (i < c.Ncap || rtag[t] == 0xffffffffu)) rtag[t] = i;
    }

cudaError_t launch_mg_rtag(const MgConst& c, const MgLocal& l, const float4* pos, unsigned int* rtag, unsigned int ntags,
                           const unsigned int* enable, cudaStream_t stream)
    {
    k_mg_rtag_clear<<<(ntags + 255u) / 256u, 256, 0, stream>>>(rtag, ntags, enable);
    k_mg_rtag_fill<<<(c.Ncap + c.Gcap + 255u) / 256u, 256, 0, stream>>>(c, l, pos, rtag, ntags, enable);
    return cudaGetLastError();
    }

// ---- launchers ------------------------------------------------------------------------------------------------------------
cudaError_t launch_mg_step_barrier(const MgConst& c, const MgLocal& l, unsigned long long cond_handle, int use_cond,
                                   cudaStream_t stream, unsigned long long* nvt_acc)
    {
    k_mg_step_barrier<<<1, 32, 0, stream>>>(c, l.ms, l.st, (cudaGraphConditionalHandle)cond_handle, use_cond, nvt_acc);
    return cudaGetLastError();
    }

cudaError_t launch_mg_barrier(const MgConst& c, const MgLocal& l, int kind, const unsigned int* enable,
                              cudaStream_t stream)
    {
    k_mg_barrier<<<1, 32, 0, stream>>>(c, l.ms, kind, enable);
    return cudaGetLastError();
    }

cudaError_t launch_mg_rebuild_vote(const MgConst& c, const MgLocal& l, int buffer, const unsigned int* enable,
                                   cudaStream_t stream)
    {
    k_mg_rebuild_vote<<<1, 32, 0, stream>>>(c, l.ms, l.st, buffer, enable);
    return cudaGetLastError();
    }

cudaError_t launch_mg_migrate(const MgConst& c, const MgLocal& l, float4* pos, const unsigned int* enable,
                              cudaStream_t stream)
    {
    const unsigned int nb = (std::max(std::max(c.Gcap, l.nchunks), (unsigned int)MG_NDIR) + 255u) / 256u;
    k_mg_begin<<<nb, 256, 0, stream>>>(c, l.ms, pos, l.chunk_flag, l.nchunks, enable);
    k_mg_migrate_mark<<<(c.Ncap + 255u) / 256u, 256, 0, stream>>>(c, l, pos, enable);
    k_mg_migrate_publish<<<1, 32, 0, stream>>>(c, l.ms, enable);
    k_mg_migrate_count<<<1, 32, 0, stream>>>(c, l, enable);
    k_mg_migrate_unpack<<<dim3((c.mig_cap + 127u) / 128u, MG_NDIR), 128, 0, stream>>>(c, l, pos, enable);
    k_mg_low_holes<<<32, 256, 0, stream>>>(l, enable);
    k_mg_move_high<<<32, 256, 0, stream>>>(l, pos, enable);
    k_mg_commit<<<1, 1, 0, stream>>>(l, enable);
    return cudaGetLastError();
    }

cudaError_t launch_mg_ghost_send(const MgConst& c, const MgLocal& l, const float4* pos, int buffer,
                                 const unsigned int* enable, cudaStream_t stream)
    {
    if (c.Gcap == 0) return cudaSuccess;
    k_mg_ghost_send<<<(c.Gcap + 255u) / 256u, 256, 0, stream>>>(c, l, pos, l.slot_dir, buffer, enable);
    return cudaGetLastError();
    }

cudaError_t launch_mg_ghost_plan(const MgConst& c, const MgLocal& l, float4* pos, int buffer, const unsigned int* enable,
                                 cudaStream_t stream)
    {
    k_mg_ghost_plan<<<(c.Ncap + 255u) / 256u, 256, 0, stream>>>(c, l, pos, enable);
    k_mg_cta_order<<<1, 1024, 0, stream>>>(l, enable);
    cudaError_t e = launch_mg_ghost_send(c, l, pos, buffer, enable, stream);
    if (e != cudaSuccess) return e;
    return launch_mg_barrier(c, l, 2, enable, stream);
    }
} // namespace b200
