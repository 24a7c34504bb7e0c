// Multi-GPU (spatial domain decomposition) device-side state and kernel launchers.  Replaces the device half of
// hoomd/CommunicatorGPU.{cc,cu} (migrateParticles :1495, exchangeGhosts :1804, beginUpdateGhosts/finishUpdateGhosts
// :2500/:2870) and hoomd/DomainDecomposition.cc:531-557 (local box) with a design for one NVSwitch box:
//   * every rank maps every other rank's "slab" (positions A/B, tags, migration mailboxes, flag words) with CUDA IPC;
//   * ghost positions are WRITTEN BY THE OWNER straight into the neighbor's position buffer over NVLink (no pack /
//     send / recv / unpack), 26 directions at once, each with its periodic shift;
//   * cross-GPU ordering uses flag words in the peer's memory (one all-ranks barrier per MD step, which also carries
//     the distance-check vote that the reference all-reduces, hoomd/md/NeighborListGPU.cc:117-122);
//   * migration / ghost-plan construction run on the device inside the same conditional rebuild bracket as the
//     single-GPU path: no host round trip per step or per rebuild.
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace b200
{
constexpr int MG_MAX_RANKS = 8;
constexpr int MG_NDIR = 27;      // direction code = (dz+1)*9 + (dy+1)*3 + (dx+1); 13 = stay
constexpr int MG_DIR_STAY = 13;
constexpr int MG_NBARRIER = 4;   // flag kinds: 0 = MD step (+ distance-check vote), 1..3 = rebuild phases

// pdata_element (hoomd/ParticleData.h:241-257) restricted to what the path carries; 64 B
struct MigElem
    {
    float4 pos;
    float4 vel;
    float ax, ay, az;
    int ix, iy, iz;
    unsigned int tag;
    unsigned int _pad;
    };

// Device-resident, per-rank mutable communication state (lives in ordinary device memory of the owner)
struct MgState
    {
    unsigned long long epoch[MG_NBARRIER];   // barriers passed so far, per kind
    unsigned int send_cnt[MG_NDIR];          // migrants packed per direction (this rebuild)
    unsigned int gsend_cnt[MG_NDIR];         // ghosts planned per direction (valid until the next rebuild)
    unsigned int recv_off[MG_NDIR + 1];      // exclusive prefix of the received migrant counts
    unsigned int n_leave;                    // holes left by migrants
    unsigned int n_low_holes;                // holes below the new N that still need a particle
    unsigned int n_moved_down;               // tail particles moved into low holes so far
    unsigned int N_new;                      // local particle count after this migration
    unsigned int rebuild_id;                 // incremented per rebuild (stamps hole marks)
    unsigned int error;                      // sticky: 1 migrate overflow, 2 ghost overflow, 4 local capacity overflow
    unsigned int max_send, max_gsend, max_N; // high-water marks for capacity tuning
    unsigned int n_bchunks;                  // particle chunks of k_md_step that touch the ghost layer (scheduled first)
    unsigned long long wait_ns[MG_NBARRIER]; // time spent between arriving at a barrier and seeing every rank, per kind
    unsigned long long wait_cnt[MG_NBARRIER];
    };

// By-value kernel argument: geometry + peer pointers (constant between reconnects)
struct MgConst
    {
    unsigned int world, rank;
    unsigned int Ncap;                       // local capacity; ghosts live at [Ncap, Ncap + Gcap)
    unsigned int Gcap;
    unsigned int mig_cap;                    // migrants per direction
    float lo[3], hi[3];                      // owned region (local box along split axes, global box otherwise)
    float Lg[3];                             // global box lengths
    float glo[3], ghi[3];                    // global box bounds
    unsigned char split[3];                  // axis is divided among > 1 ranks
    unsigned char _pad;
    float r_ghost;                           // ghost layer width = max r_cut + r_buff (NeighborList::getGhostLayerWidth)
    unsigned int seg_off[MG_NDIR];           // receiver-side ghost segment of direction d (identical on every rank)
    unsigned int seg_cap[MG_NDIR];
    float shift[MG_NDIR][3];                 // added to a position sent in direction d (periodic image of the receiver)
    int dir_rank[MG_NDIR];                   // rank of the neighbor in direction d (-1: unused direction)
    // peer views by direction (the neighbor in direction d)
    float4* peer_posA[MG_NDIR];
    float4* peer_posB[MG_NDIR];
    unsigned int* peer_tag[MG_NDIR];
    MigElem* peer_mig[MG_NDIR];              // the neighbor's receive segment for MY direction d
    unsigned int* peer_mig_cnt[MG_NDIR];     // the neighbor's receive count word for MY direction d
    // flag arrays by rank: flags[(kind*2 + parity) * MG_MAX_RANKS + writer_rank]
    unsigned long long* flags[MG_MAX_RANKS];
    // own views
    MigElem* my_mig;                         // [MG_NDIR * mig_cap], segment d holds what arrived from direction d's sender
    unsigned int* my_mig_cnt;                // [MG_NDIR]
    };

struct MgLocal // plain device arrays of the owner used by the rebuild kernels
    {
    MgState* ms;
    StepState* st;
    float4* vel; float* accel; int* image; unsigned int* tag;
    unsigned int* holes;       // [Ncap] indices vacated by migrants
    unsigned int* low_holes;   // [Ncap]
    unsigned int* hole_mark;   // [Ncap] == rebuild_id where the slot was vacated in this rebuild
    unsigned int* plan_src;    // [sum seg_cap] local index of the particle feeding each remote ghost slot
    const unsigned char* slot_dir; // [sum seg_cap] direction code of each ghost slot (constant)
    unsigned int* pplan;       // [7 * Ncap] reverse plan: ghost slots fed by particle i (k_md_step's fused ghost update)
    unsigned char* pplan_n;    // [Ncap]
    unsigned int* chunk_flag;  // [nchunks] chunk of `ppb` consecutive particles holds one within the ghost layer of a split face
    unsigned int* cta_order;   // [nchunks] boundary chunks first (ascending), then the others (ascending)
    unsigned int ppb, nchunks; // particles per CTA of k_md_step, ceil(Ncap / ppb)
    };

__device__ __forceinline__ unsigned long long globaltimer_ns()
    {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
    }

__device__ __forceinline__ void st_flag(unsigned long long* p, unsigned long long v)
    {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
    }

__device__ __forceinline__ unsigned long long ld_flag(const unsigned long long* p)
    {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
    }

// All-ranks barrier, executed by ONE warp (all 32 lanes must call it).  Flag word = (epoch << 1) | payload bit.  Two slots per
// kind (epoch parity): a fast rank may already publish epoch+1 while a slow one still reads epoch.  Returns the OR of the
// payload bits.  A peer that never arrives (crashed process) must not hang the GPU: after 4 s the wait gives up and sets the
// sticky error bit 8; the host reports it at the end of the chunk.  flags[q] = rank q's flag array,
// flags[q][(kind*2 + parity) * MG_MAX_RANKS + writer].
// value / sum_out (optional): an all-reduce riding on the barrier -- every rank deposits one 64-bit integer in every rank's
// value slots (behind the flag words, same indexing) before its flag, and adds up the world's deposits after the wait.  Integer
// addition: every rank gets the same bits (TwoStepNVTMTK's kinetic energy under decomposition, fixed point).
__device__ __forceinline__ bool mg_warp_barrier(unsigned int world, unsigned int rank, unsigned long long* const* flags,
                                                MgState* ms, int kind, bool payload, unsigned long long value = 0ull,
                                                unsigned long long* sum_out = nullptr)
    {
    const unsigned int lane = threadIdx.x & 31;
    const unsigned long long epoch = ms->epoch[kind] + 1ull;
    const unsigned int slot = ((unsigned int)kind * 2u + (unsigned int)(epoch & 1ull)) * MG_MAX_RANKS;
    constexpr unsigned int VAL0 = MG_NBARRIER * 2u * MG_MAX_RANKS; // value slots follow the flag words
    if (sum_out && lane < world) st_flag(flags[lane] + VAL0 + slot + rank, value);
    __threadfence_system(); // everything this rank wrote (ghosts, migrants, counts) is visible before the flag is
    __syncwarp();
    const unsigned long long t0 = globaltimer_ns();
    if (lane < world) st_flag(flags[lane] + slot + rank, (epoch << 1) | (payload ? 1ull : 0ull));
    bool bit = false;
    if (lane < world)
        {
        const unsigned long long* mine = flags[rank] + slot + lane;
        unsigned long long v = ld_flag(mine);
        while ((v >> 1) < epoch)
            {
            if (globaltimer_ns() - t0 > 4000000000ull)
                {
                atomicOr(&ms->error, 8u);
                break;
                }
            v = ld_flag(mine);
            }
        bit = (v & 1ull) != 0ull;
        }
    const bool any = __any_sync(0xffffffffu, bit);
    __threadfence_system();
    __syncwarp();
    if (sum_out)
        {
        unsigned long long v = lane < world ? ld_flag(flags[rank] + VAL0 + slot + lane) : 0ull;
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        *sum_out = v;
        }
    if (lane == 0)
        {
        ms->epoch[kind] = epoch;
        ms->wait_ns[kind] += globaltimer_ns() - t0;
        ms->wait_cnt[kind] += 1ull;
        }
    __syncwarp();
    return any;
    }

// all-ranks barrier of kind k; kind 0 also carries the distance-check vote and takes the rebuild decision
// nvt_acc != NULL (TwoStepNVTMTK): the barrier also sums the ranks' kinetic-energy accumulators of the step about to run
cudaError_t launch_mg_step_barrier(const MgConst& c, const MgLocal& l, unsigned long long cond_handle, int use_cond,
                                   cudaStream_t stream, unsigned long long* nvt_acc = nullptr);
cudaError_t launch_mg_barrier(const MgConst& c, const MgLocal& l, int kind, const unsigned int* enable,
                              cudaStream_t stream);
// rebuild bracket, last: collective Nmax-overflow vote (sets st->abort on every rank)
cudaError_t launch_mg_rebuild_vote(const MgConst& c, const MgLocal& l, int buffer, const unsigned int* enable,
                                   cudaStream_t stream);
// rebuild bracket, part 1: migration (mark + pack into the neighbor's mailbox) ... barrier ... unpack + hole filling
cudaError_t launch_mg_migrate(const MgConst& c, const MgLocal& l, float4* pos, const unsigned int* enable,
                              cudaStream_t stream);
// rebuild bracket, part 2: ghost plan + first ghost write (+ tags) ... barrier
cudaError_t launch_mg_ghost_plan(const MgConst& c, const MgLocal& l, float4* pos, int buffer, const unsigned int* enable,
                                 cudaStream_t stream);
// rebuild bracket, after the ghost plan: tag -> index over locals and ghosts (bonds / exclusions keyed by global tag)
cudaError_t launch_mg_rtag(const MgConst& c, const MgLocal& l, const float4* pos, unsigned int* rtag, unsigned int ntags,
                           const unsigned int* enable, cudaStream_t stream);
// every step: owners write their boundary particles into the neighbors' ghost slots of buffer `buffer` (0 = A, 1 = B)
cudaError_t launch_mg_ghost_send(const MgConst& c, const MgLocal& l, const float4* pos, int buffer,
                                 const unsigned int* enable, cudaStream_t stream);
} // namespace b200
