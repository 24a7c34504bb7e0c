// Shared pieces of the single-kernel MD step: the argument block, the per-particle integration half (step two of the
// step in flight + step one of the next + distance check) and the last-CTA rebuild decision.  Used by the CSR-walk
// kernel (fused_step.cu) and the shared-memory cell-tile kernel (tile.cu) => identical arithmetic per particle.
#pragma once
#include "common.cuh"
#include <cstring>
#include "integrate_body.cuh"
#include "pair_body.cuh"
#include "kernels.h"
#include "decide.cuh"
#include "tile.h"
#include "comm.cuh"

namespace b200
{
// ---- the single-kernel MD step -------------------------------------------------------------------------------------------
// One launch = PotentialPair::computeForces(ts+1) + integrateStepTwo(ts) + integrateStepOne(ts+1) + distanceCheck +
// NeighborList::needsUpdating(ts+2).  Velocity Verlet lets step two of one update() and step one of the next run back
// to back per particle (nothing in between reads another particle), so the only grid-wide dependencies left in a
// step are "all positions written before any force reads them" -- solved by ping-pong position buffers (pos_in is read
// by every neighbor gather, pos_out is written by the owner) -- and the rebuild decision, taken by the last CTA to
// finish (ticket counter).  Arithmetic per particle is the shared bodies => bit-identical to the separate kernels.
template<class Evaluator> struct MdStepArgs
    {
    const float4* pos_in; float4* pos_out; float4* vel; float3s* accel; int3* image; float4* net_force;
    float* virial; unsigned int virial_pitch;
    const float4* last_pos; const unsigned int* tag;
    const unsigned int* n_neigh; const unsigned int* nlist; const unsigned int* head_list;
    const typename Evaluator::param_type* params; const float* rcutsq; const float* ronsq; unsigned int ntypes;
    const float* maxsq; const float* gamma;
    unsigned int N; Box box; float dt; float r_buff; float wrap_skip_margin;
    int method; int limit; float limit_val; int zero_force; float kT; unsigned int seed; int noiseless_t;
    int copy_pos_on_last;     // this launch reads the alternate buffer: the final positions must land in pos_out
    const unsigned int* d_N;  // multi-GPU: live local particle count (N is the capacity the grid was sized for)
    int mg;                   // multi-GPU: 1 = the rebuild decision is left to the stand-alone step-barrier kernel; 2 = the
                              // LAST CTA of this launch takes the all-ranks barrier (vote + decision) itself
    StepState* st; cudaGraphConditionalHandle handle; int use_cond;
    // cell-tile path (tile.cu)
    TileGeom tile; const unsigned int* cell_start; const uint4* ell4; unsigned int ell_pitch, ell_nmax;
    int force_only;           // compute and store the net force only (prepRun / stand-alone force evaluation)
    const float4* extra_force; const float* extra_virial; unsigned int extra_virial_pitch; // second ForceCompute (bonds)
    GhostSendArgs gs; int gs_buffer; // multi-GPU: fused ghost update into the peers' buffer gs_buffer (0 = A, 1 = B)
    NvtState* nvt; unsigned long long* nvt_acc; double nvt_scale; unsigned int nvt_ndof; float nvt_tau; // TwoStepNVTMTK
    const unsigned char* group_flag;                                  // the method's group by tag (NULL = all)
    const float* diameter; int use_lambda; float gamma_lambda; double* tally_acc; // langevin dscale / tally
    unsigned int nmax_uniform; int prefetch;                          // see FusedStepArgs
    int persist_strict;                                               // k_md_steps: acquire fence behind the spin (md_steps_persist.cuh)
    unsigned int persist_sleep_ns;                                    // k_md_steps: pause between two polls of the release word
    };

__device__ __forceinline__ void prefetch_l1(const void* p)
    {
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
    }

// write particle i's position into every ghost slot it feeds on the neighbor ranks (plain stores over NVLink)
template<class Evaluator>
__device__ __forceinline__ bool md_step_send_ghosts(const MdStepArgs<Evaluator>& a, unsigned int i, const float4 p)
    {
    const unsigned int n = a.gs.pplan_n[i];
    for (unsigned int k = 0; k < n; k++)
        {
        const unsigned int e = a.gs.pplan[(size_t)k * a.gs.pitch + i]; // direction in the top 5 bits, slot below
        const unsigned int t = e & 0x07ffffffu, d = e >> 27;
        float4* dst = a.gs_buffer ? a.gs.tab->peer_posB[d] : a.gs.tab->peer_posA[d];
        dst[a.gs.pitch + t] = p; // raw coordinates (see k_mg_ghost_send)
        }
    return n != 0u;
    }

#ifndef B200_MD_BLOCK
#define B200_MD_BLOCK 128
#endif
constexpr int MD_BLOCK = B200_MD_BLOCK; // threads per CTA of k_md_step (B200 sweep: 128 beats 256 by 1.2 % at 64k and 1M)

// phase 2 of the step for particle i: net force F -> step two(ts2) -> step one(ts2+1) -> distance check.  Returns the
// particle's distance-check result.
// EXTRAS = false compiles out the integration group, the Langevin dscale / tally options and the multi-GPU ghost stores:
// the plain single-GPU variants keep their 48-register budget without spills (md_step_kernel.cuh)
template<class Evaluator, bool VIRIAL, bool EXTRAS = true>
__device__ __forceinline__ bool md_step_integrate(const MdStepArgs<Evaluator>& a, unsigned int i, float4 F,
                                                  const float4 pi4, float (&vir)[6], unsigned int ts2, bool last,
                                                  bool store_force, double& mv2, float exp_fac = 1.0f,
                                                  bool* pushed = nullptr)
    {
    bool moved = false;
    float4 v = a.vel[i];
    if (a.extra_force)
        {
        // Integrator::computeNetForce: net = pair + bond, in the order the forces were added (hoomd/Integrator.cc:308-420)
        const float4 e = a.extra_force[i];
        F.x = __fadd_rn(F.x, e.x); F.y = __fadd_rn(F.y, e.y); F.z = __fadd_rn(F.z, e.z); F.w = __fadd_rn(F.w, e.w);
        if (VIRIAL)
            {
#pragma unroll
            for (int c = 0; c < 6; c++) vir[c] = __fadd_rn(vir[c], a.extra_virial[(size_t)c * a.extra_virial_pitch + i]);
            }
        }
    if (store_force)
        {
        a.net_force[i] = F;
        if (VIRIAL)
            {
#pragma unroll
            for (int c = 0; c < 6; c++) a.virial[c * a.virial_pitch + i] = vir[c];
            }
        }
    const unsigned int type_i = (unsigned int)__float_as_int(pi4.w);
    if (EXTRAS && a.group_flag && !a.group_flag[a.tag[i]])
        {
        // not in the integration method's group: the force is reported, nothing moves (the ping-pong buffer still needs
        // the unchanged position; a particle at rest never trips the distance check)
        if (!last || a.copy_pos_on_last) a.pos_out[i] = pi4;
        return false;
        }
    const bool nve = a.method == 1, nvt = a.method == 3;
    // exp_fac: ONE thermostat factor serves step two(ts) and step one(ts+1) -- advanceThermostat runs between step one and
    // step two of the same update() (hoomd/md/TwoStepNVTMTK.cc:296), i.e. between two launches (md_step_nvt_begin)
    float3s acc;
    // ---- integrateStepTwo(ts) ----
    if (nve)
        nve_step_two_body(v, acc, F, a.dt, a.limit != 0, a.limit_val, a.zero_force != 0);
    else if (nvt)
        nvt_step_two_body(v, acc, F, a.dt, exp_fac);
    else
        {
        const unsigned int ptag = a.tag[i];
        const float gamma = (EXTRAS && a.use_lambda) ? __fmul_rn(a.gamma_lambda, a.diameter[ptag]) : a.gamma[type_i];
        const float e = langevin_step_two_body(v, acc, F, ptag, ts2, a.seed, gamma, a.kT, a.dt, a.noiseless_t != 0, 3u);
        if (EXTRAS && a.tally_acc) mv2 = (double)e; // energy handed to the particle by the bath this step (md_step_finish sums it)
        }
    a.accel[i] = acc;
    if (!last)
        {
        // ---- integrateStepOne(ts+1): Langevin step one == NVE step one without limit (TwoStepLangevin.cc:86-125) ----
        float4 p = pi4;
        int3 img = a.image[i];
        const float4 lp = a.last_pos[i];
        if (nvt)
            {
            nvt_step_one_body(p, v, acc, img, a.box, a.dt, exp_fac);
            mv2 = nvt_mv2(v); // summand of the kinetic energy advanceThermostat needs (md_step_finish reduces it)
            }
        else
            nve_step_one_body(p, v, acc, img, a.box, a.dt, nve && a.limit, a.limit_val);
        a.pos_out[i] = p;
        if (EXTRAS && a.gs.tab)
            {
            const bool sent = md_step_send_ghosts(a, i, p);
            if (pushed) *pushed = sent;
            }
        a.image[i] = img;
        // ---- NeighborList::distanceCheck (hoomd/md/NeighborList.cc:981-1016): row 0 of the table for a fixed box
        // (lambda = 1), row 1 while the box differs from the one the list was built in ----
        const float* tab = a.maxsq + (a.st->box_scaled ? a.ntypes + 3u : 0u);
        const float maxsq = tab[type_i];
        float3 dx = make_float3(__fsub_rn(p.x, __fmul_rn(tab[a.ntypes], lp.x)),
                                __fsub_rn(p.y, __fmul_rn(tab[a.ntypes + 1u], lp.y)),
                                __fsub_rn(p.z, __fmul_rn(tab[a.ntypes + 2u], lp.z)));
        dx = min_image_exact_sel(dx, a.box, a.box.periodic.x, a.box.periodic.y, a.box.periodic.z);
        moved = dot_exact(dx) >= maxsq;
        }
    else if (a.copy_pos_on_last)
        {
        a.pos_out[i] = pi4;
        if (EXTRAS && a.gs.tab)
            {
            const bool sent = md_step_send_ghosts(a, i, pi4);
            if (pushed) *pushed = sent;
            }
        }
    a.vel[i] = v;
    return moved;
    }

// ---- TwoStepNVTMTK inside the step kernel ---------------------------------------------------------------------------------
// The thermostat needs sum(m v^2) of the velocities step one left behind.  Every CTA adds its share, converted to fixed
// point (a power-of-two scale: exact conversion of the double up to its rounding to an integer), into one of NVT_SLOTS
// 64-bit accumulators with a fire-and-forget atomic: integer addition is associative, so the total is the same bits
// whatever the order -- deterministic without partial-sum arrays, fences or a serial tail.  The NEXT launch's first thread
// (block 0; the kernel boundary orders it after all the adds) turns the total into xi / eta / exp_fac and publishes them with
// the launch's timestep as tag; the other CTAs meet the tag only after their force phase, microseconds later (bounded spin).
template<class Evaluator>
__device__ __forceinline__ void md_step_nvt_begin(const MdStepArgs<Evaluator>& a, StepState* st, unsigned int ts2)
    {
    if (a.method != 3 || blockIdx.x != 0 || threadIdx.x >= 32) return; // warp 0 of block 0, one accumulator per lane
    if (((volatile NvtState*)a.nvt)->gen == ts2) return; // the chunk prologue advanced the thermostat for this step
    unsigned long long* acc = a.nvt_acc + (size_t)(ts2 & 1u) * NVT_SLOTS * NVT_SLOT_STRIDE;
    unsigned long long v = 0ull;
    if (threadIdx.x < NVT_SLOTS)
        {
        v = __ldcg(acc + threadIdx.x * NVT_SLOT_STRIDE);
        acc[threadIdx.x * NVT_SLOT_STRIDE] = 0ull; // ready for the launch after the next one
        }
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0)
        {
        nvt_advance_body(a.nvt, (double)(long long)v / a.nvt_scale, a.nvt_ndof, a.kT, a.nvt_tau, a.dt);
        __threadfence();
        ((volatile NvtState*)a.nvt)->gen = ts2;
        }
    (void)st;
    }

// thread 0 of every CTA, before the barrier between the force phase and the integration phase: wait for this launch's
// thermostat factor and hand it to the CTA through shared memory
template<class Evaluator>
__device__ __forceinline__ void md_step_nvt_wait(const MdStepArgs<Evaluator>& a, StepState* st, unsigned int ts2,
                                                 float* s_exp_fac)
    {
    if (a.method != 3 || threadIdx.x != 0) return;
    volatile NvtState* nvt = (volatile NvtState*)a.nvt;
    unsigned int spins = 0;
    while (nvt->gen != ts2)
        {
        __nanosleep(64);
        if (++spins > (1u << 21)) // ~0.1 s: block 0 never ran ahead of this CTA; fail loudly instead of hanging
            {
            st->nvt_error = 2u;
            break;
            }
        }
    __threadfence();
    *s_exp_fac = nvt->exp_fac;
    }

// last CTA to finish takes the rebuild decision for the next launch (every thread of the CTA must call this).
template<class Evaluator, bool EXTRAS = true>
__device__ __forceinline__ void md_step_finish(const MdStepArgs<Evaluator>& a, StepState* st, bool moved,
                                               unsigned int steps_left, bool last, double mv2 = 0.0,
                                               unsigned int ts2 = 0, bool pushed = false)
    {
    const int any_moved = __syncthreads_or(moved ? 1 : 0);
    // multi-GPU: did this CTA store ghosts into a peer's buffer?  Only those CTAs (the boundary chunks, scheduled first) pay
    // for a system-scope fence; a fence in every CTA cost 15-20 % of the kernel (profiles/r2e_*)
    const int any_pushed = (EXTRAS && a.mg == 2) ? __syncthreads_or(pushed ? 1 : 0) : 0;
    if (EXTRAS && a.method == 2 && a.tally_acc)
        {
        // reservoir-energy tally: sum_j F_bd . v' of this step (double atomics: a diagnostic, order-dependent in the last bits)
        for (int o = 16; o > 0; o >>= 1) mv2 += __shfl_xor_sync(0xffffffffu, mv2, o);
        if ((threadIdx.x & 31) == 0 && mv2 != 0.0)
            {
            atomicAdd(a.tally_acc, mv2);
            atomicAdd(a.tally_acc + 1u + (ts2 & 1u), mv2);
            }
        if (blockIdx.x == 0 && threadIdx.x == 0) a.tally_acc[1u + ((ts2 + 1u) & 1u)] = 0.0; // slot of the next step
        }
    if (a.method == 3 && !last)
        {
        // this CTA's share of sum(m v^2): warp shuffles, then one fixed-point atomic per warp that integrated particles
        for (int o = 16; o > 0; o >>= 1) mv2 += __shfl_xor_sync(0xffffffffu, mv2, o);
        if ((threadIdx.x & 31) == 0 && mv2 != 0.0)
            {
            const double scaled = mv2 * a.nvt_scale;
            // 2^62 spread over every warp of the grid: no sum of admissible shares can wrap the accumulator
            const double limit = 4611686018427387904.0 / ((double)gridDim.x * (MD_BLOCK / 32));
            if (!(scaled < limit))
                st->nvt_error = 1u;
            else
                {
                const unsigned int slot = (blockIdx.x * (MD_BLOCK / 32) + (threadIdx.x >> 5)) % NVT_SLOTS;
                atomicAdd(a.nvt_acc + ((size_t)((ts2 + 1u) & 1u) * NVT_SLOTS + slot) * NVT_SLOT_STRIDE,
                          (unsigned long long)__double2ll_rn(scaled));
                }
            }
        }
    __shared__ int s_is_last;
    if (threadIdx.x == 0)
        {
        // one 64-bit atomic carries both the ticket and this CTA's distance-check result: no fence, no second atomic
        const unsigned long long inc = 1ull | (any_moved ? (1ull << 32) : 0ull);
        if (any_pushed) __threadfence_system(); // this CTA's ghost stores (ordered by the barrier above) before its ticket
        const unsigned long long old = atomicAdd(&st->ticket, inc);
        const bool is_last = (unsigned int)old == gridDim.x * gridDim.y * gridDim.z - 1;
        s_is_last = is_last ? 1 : 0;
        if (is_last && !(EXTRAS && a.mg))
            {
            // One GPU: the rebuild decision on a register copy of the control block (four independent 16-byte loads, one round
            // trip, then four stores) instead of a dozen dependent loads at the tail of every launch; nobody else touches
            // the block once every CTA has taken its ticket.  Same decision, same fields written.
            static_assert(sizeof(StepState) == 96, "StepState is copied as six 16-byte words");
            StepState loc;
            uint4* lq = reinterpret_cast<uint4*>(&loc);
            const uint4* gq = reinterpret_cast<const uint4*>(st);
#pragma unroll
            for (int q = 0; q < 6; q++) lq[q] = __ldcg(gq + q);
            loc.moved = ((old + inc) >> 32) ? 1u : 0u;
            loc.ticket = 0ull;
            loc.steps_left = steps_left - 1u;
            if (!last)
                decide_body<false>(&loc);
            else
                {
                loc.need_rebuild = 0u;
                loc.moved = 0u;
                }
            uint4* wq = reinterpret_cast<uint4*>(st);
            wq[0] = lq[0]; // timestep, ts_step_two, moved, need_rebuild
            wq[1] = lq[1]; // force_update, last_updated, every, r_buff_tiny
            wq[2] = lq[2]; // num_updates, num_forced, overflow, ntypes
            wq[3] = lq[3]; // steps_left, epoch, ticket
            st->box_scaled = loc.box_scaled;
#if (CUDART_VERSION >= 12040)
            if (a.use_cond) cudaGraphSetConditional(a.handle, loc.need_rebuild);
#endif
            }
        else if (is_last)
            {
            if ((old + inc) >> 32) st->moved = 1u;
            st->ticket = 0ull;
            st->steps_left = steps_left - 1u;
            // the vote needs every rank: the step barrier decides (below, or k_mg_step_barrier)
#if (CUDART_VERSION >= 12040)
            if (a.use_cond && a.mg != 2) cudaGraphSetConditional(a.handle, st->need_rebuild);
#endif
            }
        }
    if (EXTRAS && a.mg == 2)
        {
        // Multi-GPU: the last CTA of this rank to finish takes the all-ranks step barrier.  Every CTA's ghost stores into the
        // peers' buffers are fenced before its ticket, so the flag this warp publishes orders them (RAW for the neighbors'
        // next force phase); the wait orders every rank's reads of this step before anybody's next ghost stores (WAR); the
        // payload bit is the global distance-check vote (hoomd/md/NeighborListGPU.cc:117-122), after which every rank takes
        // the same NeighborList::needsUpdating decision.  The boundary chunks ran FIRST (cta_order), so their NVLink stores
        // have long landed when the barrier is reached: the exchange overlaps the interior force work.
        __syncthreads();
        if (s_is_last && threadIdx.x < 32)
            {
            const GhostSendTable* t = a.gs.tab;
            const bool moved_any = mg_warp_barrier(t->world, t->rank, t->flags, t->ms, 0,
                                                   !last && *((volatile unsigned int*)&st->moved) != 0u);
            if (threadIdx.x == 0)
                {
                st->moved = moved_any ? 1u : 0u;
                if (!last && !st->abort)
                    decide_body(st);
                else
                    {
                    st->need_rebuild = 0u;
                    st->moved = 0u;
                    }
#if (CUDART_VERSION >= 12040)
                if (a.use_cond) cudaGraphSetConditional(a.handle, st->need_rebuild);
#endif
                }
            }
        }
    }
// host: the argument block from the engine's descriptors (shared by the CSR-walk and the cell-tile launchers)
template<class Evaluator>
inline void fill_md_step_args(MdStepArgs<Evaluator>& a, const b200_pair_args& pa, const float* d_params,
                              const FusedStepArgs& fa, const float4* pos_in, float4* pos_out, int copy_pos_on_last,
                              StepState* st, unsigned long long cond_handle, int use_cond)
    {
    memset(&a, 0, sizeof(a));
    a.pos_in = pos_in; a.pos_out = pos_out; a.vel = fa.vel; a.accel = (float3s*)fa.accel; a.image = (int3*)fa.image;
    a.net_force = (float4*)pa.d_force; a.virial = pa.d_virial; a.virial_pitch = pa.virial_pitch;
    a.last_pos = fa.last_pos; a.tag = fa.tag;
    a.n_neigh = pa.d_n_neigh; a.nlist = pa.d_nlist; a.head_list = pa.d_head_list;
    a.params = (const typename Evaluator::param_type*)d_params; a.rcutsq = pa.d_rcutsq; a.ronsq = pa.d_ronsq;
    a.ntypes = pa.ntypes;
    a.maxsq = fa.maxsq; a.gamma = fa.gamma;
    // fa.box locates and wraps the particles (the local box); pa.box is the box minimum images are taken in
    a.N = fa.N; a.box = make_box_image(fa.box, &pa.box); a.dt = fa.dt; a.r_buff = fa.r_buff;
    a.wrap_skip_margin = pa.wrap_skip_margin;
    a.method = fa.method; a.limit = fa.limit; a.limit_val = fa.limit_val; a.zero_force = fa.zero_force;
    a.kT = fa.kT; a.seed = fa.seed; a.noiseless_t = fa.noiseless_t;
    a.copy_pos_on_last = copy_pos_on_last;
    a.d_N = fa.d_N; a.mg = fa.mg;
    a.extra_force = fa.extra_force; a.extra_virial = fa.extra_virial; a.extra_virial_pitch = fa.extra_virial_pitch;
    a.gs = fa.gs; a.gs_buffer = copy_pos_on_last ? 0 : 1; // phase 0 writes B, phase 1 writes A
    a.nvt = (NvtState*)fa.nvt_state; a.nvt_acc = fa.nvt_acc; a.nvt_scale = fa.nvt_scale; a.nvt_ndof = fa.nvt_ndof;
    a.nvt_tau = fa.nvt_tau;
    a.group_flag = fa.group_flag;
    a.diameter = fa.diameter; a.use_lambda = fa.use_lambda; a.gamma_lambda = fa.gamma_lambda; a.tally_acc = fa.tally_acc;
    a.nmax_uniform = fa.nmax_uniform; a.prefetch = fa.prefetch;
    a.st = st; a.handle = (cudaGraphConditionalHandle)cond_handle; a.use_cond = use_cond;
    }
} // namespace b200
