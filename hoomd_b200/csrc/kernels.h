// Internal (C++) declarations shared between the kernel translation units and the engine.  Not part of the C ABI.
#pragma once
#include <cstddef>
#include <cuda_runtime.h>
#include "../../include/hoomd_b200.h"

namespace b200
{
// Device-resident control block of the device-driven step loop (one per system).  Everything the reference keeps in
// host members of NeighborList (m_last_updated_tstep, m_force_update, m_updates, ...; hoomd/md/NeighborList.cc:1080-1163)
// lives here so that no step needs a host round trip.
struct StepState
    {
    unsigned int timestep;        // System::m_cur_tstep
    unsigned int ts_step_two;     // the `timestep` argument of integrateStepTwo of the step in flight
    unsigned int moved;           // distance-check result of the step in flight (set by the step-one kernel)
    unsigned int need_rebuild;    // consumed by the rebuild kernels (`enable` pointer) / the conditional graph node
    unsigned int force_update;    // NeighborList::m_force_update
    unsigned int last_updated;    // NeighborList::m_last_updated_tstep
    unsigned int every;           // NeighborList::m_every (check_period)
    unsigned int r_buff_tiny;     // r_buff < 1e-6: rebuild whenever allowed
    unsigned int num_updates;     // NeighborList::m_updates
    unsigned int num_forced;      // NeighborList::m_forced_updates
    unsigned int overflow;        // sticky: some n_neigh exceeded Nmax during a device-driven rebuild
    unsigned int ntypes;
    unsigned int steps_left;      // MD steps still to run in this chunk (k_md_step counts it down; 0 = launches are no-ops)
    unsigned int epoch;           // k_md_steps: grid-wide release word, +1 per MD step of a multi-step launch
    unsigned long long ticket;    // k_md_step: low word = CTAs finished, high word = CTAs that saw a mover (one atomic)
    unsigned int store_force;     // write net_force/virial every step (otherwise only in the last step of a chunk)
    unsigned int N;               // local particle count (multi-GPU: changes at every migration)
    unsigned int abort;           // multi-GPU: some rank's rebuild overflowed Nmax; every rank stops stepping (1 + buffer)
    unsigned int nvt_error;       // NVT: 1 = kinetic energy outside the accumulator range, 2 = thermostat hand-off timed out
    unsigned int no_dist_check;   // NeighborList::m_dist_check off: rebuild when the check period is hit (NeighborList.cc:1126-1135)
    unsigned int box_scaled;      // the box changed since the last list build: the distance check uses row 1 of the
                                  // threshold table (lambda-scaled, NeighborList.cc:981-1007); cleared by the rebuild decision
    unsigned long long ctl64;     // (reserved)
    };
// The step kernels copy the block as six 16-byte words and write back words 0-3 plus box_scaled (md_step_body.cuh,
// md_steps_persist.cuh): the fields other kernels or CTAs may write while a step kernel's last CTA does so must stay outside
// words 0-3 -- store_force, N, abort, nvt_error, no_dist_check -- and the layout below is what those copies assume.
static_assert(sizeof(StepState) == 96 && offsetof(StepState, need_rebuild) == 12 && offsetof(StepState, num_updates) == 32
                  && offsetof(StepState, overflow) == 40 && offsetof(StepState, steps_left) == 48
                  && offsetof(StepState, ticket) == 56 && offsetof(StepState, store_force) == 64
                  && offsetof(StepState, abort) == 72 && offsetof(StepState, nvt_error) == 76
                  && offsetof(StepState, box_scaled) == 84 && offsetof(StepState, ctl64) == 88,
              "StepState layout");
// The engine allocates STEP_STATE_ALLOC StepState slots: slot 0 is the state; the bytes from STEP_SYNC_OFFSET on hold the
// barrier lines of k_md_steps (md_steps_persist.cuh): its ticket on a 128-byte line of its own and STEP_SYNC_COPIES copies of
// the release word, one line each.
constexpr unsigned int STEP_STATE_ALLOC = 48, STEP_SYNC_OFFSET = 1024, STEP_SYNC_COPIES = 16, STEP_SYNC_LINE = 128;
static_assert(STEP_SYNC_OFFSET + (1 + STEP_SYNC_COPIES) * STEP_SYNC_LINE <= STEP_STATE_ALLOC * sizeof(StepState),
              "barrier lines of k_md_steps must fit behind the step state");

// d_N != NULL (multi-GPU): a->N is the ghost base (local capacity) and *d_N the live local count
cudaError_t launch_cell_list(const b200_cell_list_args* a, const unsigned int* enable, cudaStream_t stream,
                             const unsigned int* d_N = nullptr);
// launches of one launch_cell_list (have_barrier: the caller passes b200_cell_list_args.d_fused_barrier and no ghost shifts)
unsigned int cell_list_kernels(unsigned int ncells, unsigned int ntot, bool have_barrier = false);
// Device-driven stepping: the build kernel itself raises StepState::overflow when a row exceeds Nmax (what the host's
// checkConditions does after the fact, hoomd/md/NeighborList.cc:1389-1405) and, on a single GPU, zeroes steps_left so the
// rest of the chunk turns into no-ops and `conditions` keeps exactly the counts of the overflowing rebuild.
constexpr unsigned int NVT_SLOTS = 16, NVT_SLOT_STRIDE = 4; // 16 accumulators per parity, one 32 B sector each
// ComputeThermo (KE of the group = all particles) + TwoStepNVTMTK::advanceThermostat; gen != 0xffffffff tags the state
cudaError_t launch_nvt_advance(float* d_nvt_state, double* d_partial, unsigned int n_partial_blocks, const float* d_vel,
                               const unsigned int* d_group, unsigned int group_size, unsigned int ndof, float T, float tau,
                               float dt, unsigned int gen, cudaStream_t stream);

cudaError_t launch_nvt_seed(unsigned long long* d_acc, double* d_partial, unsigned int n_partial_blocks, const float* d_vel,
                            const unsigned int* d_group, unsigned int group_size, double scale, unsigned int parity,
                            cudaStream_t stream);

struct OverflowHook
    {
    unsigned int* overflow = nullptr;   // set to 1 on overflow
    unsigned int* steps_left = nullptr; // set to 0 on overflow (single GPU)
    };
cudaError_t launch_nlist_binned(const b200_nlist_args* a, const unsigned int* enable, cudaStream_t stream,
                                OverflowHook hook = OverflowHook());
unsigned int pick_pair_tpp(unsigned int N);
unsigned int md_step_block(); // threads per CTA of k_md_step (particles per CTA = md_step_block() / lanes per particle)

// Multi-GPU: ghost update fused into the MD step.  The owner of a boundary particle writes its new position (+ the
// periodic shift of the direction) straight into the neighbor ranks' position buffers over NVLink, from the same lane
// that has just integrated it -- the stores of the first CTAs overlap the force work of the CTAs still running.
// (hoomd/CommunicatorGPU.cc:2500-2870 beginUpdateGhosts / finishUpdateGhosts: pack, MPI send/recv, unpack.)
struct MgState;
struct GhostSendTable            // device-resident, constant between reconnects
    {
    float shift[27][3];
    float4* peer_posA[27];
    float4* peer_posB[27];
    // the all-ranks step barrier taken by the LAST CTA of k_md_step (comm.cuh:mg_warp_barrier): flag arrays of every rank
    unsigned int world, rank;
    unsigned long long* flags[8];
    MgState* ms;
    };
struct GhostSendArgs
    {
    const GhostSendTable* tab;   // NULL = single GPU / ghost update by the stand-alone kernel
    const unsigned int* pplan;   // [7 * pitch]: (direction << 27 | ghost slot) fed by particle i at [k * pitch + i], k < pplan_n[i]
    const unsigned char* pplan_n;
    const unsigned char* slot_dir;
    unsigned int pitch;          // local capacity Ncap == ghost base of every rank's position buffer
    // boundary-first scheduling: CTA b works on particle chunk cta_order[b]; the chunks that hold a particle within the ghost
    // layer of a split face come first, so their ghost stores cross NVLink while the interior chunks still compute
    const unsigned int* cta_order;
    };

// fused_step.cu -- kernels of the device-driven step loop (group = all particles)
struct FusedStepArgs
    {
    float4* pos; float4* vel; float* accel; int* image; const float4* net_force; const float4* last_pos;
    const unsigned int* tag; const float* rcut_max; const float* gamma;
    const float* maxsq;           // distance-check table, 2 rows of ntypes + 3 floats: threshold^2 per type, then lambda x/y/z
                                  // (NeighborList.cc:981-1006); row 0 = fixed box (lambda 1), row 1 = box scaled since the build
    float lambda[3]; float lambda_min; // the same for the chunk prologue's step one (by value)
    unsigned int N; b200_box box; float dt; float r_buff;
    int method;                   // 1 = NVE, 2 = Langevin, 3 = NVT (MTK)
    int limit; float limit_val; int zero_force;
    float kT; unsigned int seed; int noiseless_t;
    const unsigned int* d_N;      // multi-GPU: live local particle count (N above is then the capacity); else NULL
    int mg;                       // multi-GPU: the rebuild decision is taken by the all-ranks step barrier, not here
    // second ForceCompute (bond force), already evaluated at the positions this step reads: added to the pair force
    // before step two (Integrator::computeNetForce).  NULL = the pair force is the only force.
    const float4* extra_force; const float* extra_virial; unsigned int extra_virial_pitch;
    GhostSendArgs gs;             // multi-GPU: ghost update fused into k_md_step
    // TwoStepNVTMTK: thermostat state {xi, eta, exp_fac, gen} on the device; fixed-point accumulators of sum(m v^2)
    // (2 parities x NVT_SLOTS slots, NVT_SLOT_STRIDE words apart) and their scale (a power of two); NDOF of the group,
    // coupling period tau; kT above is the set point
    float* nvt_state; unsigned long long* nvt_acc; double nvt_scale; unsigned int nvt_ndof; float nvt_tau;
    // the integration method's ParticleGroup (NULL = group.all()): 1 byte per particle TAG, non-members are not integrated
    // (the reference iterates m_group->getIndexArray(), hoomd/md/TwoStepNVEGPU.cc:48-70)
    const unsigned char* group_flag;
    // md.integrate.langevin(dscale=lambda): gamma = lambda * diameter[tag]; tally: reservoir energy bookkeeping
    // (hoomd/md/TwoStepLangevin.cc:287-318,416-417): acc[0] = running sum of E_step, acc[1 + (ts & 1)] = E of step ts
    const float* diameter; int use_lambda; float gamma_lambda; double* tally_acc;
    // latency trims of k_md_step: nmax_uniform != 0 (one particle type): head_list[i] == i * Nmax[0] (the exclusive scan of
    // a constant), so the first index chunk of a row can be requested without waiting for the head list; prefetch != 0: the
    // integration phase's operands (velocity, image, last position) are pulled into L1 while the neighbor walk runs
    unsigned int nmax_uniform; int prefetch;
    };
// step one (+ distance check into st->moved)
cudaError_t launch_fused_step_one(const FusedStepArgs& a, StepState* st, cudaStream_t stream);
// rebuild decision for timestep+1 (NeighborList::needsUpdating)
cudaError_t launch_decide(StepState* st, cudaStream_t stream);
// first node of the step graph: arm the first conditional node from st->need_rebuild
cudaError_t launch_graph_pre(StepState* st, unsigned long long cond_handle, cudaStream_t stream);
// single-kernel MD step: pair force + step two + step one of the next step + distance check + rebuild decision
cudaError_t launch_md_step_lj(const b200_pair_args& pa, const float* d_params, const FusedStepArgs& fa,
                              const float4* pos_in, float4* pos_out, int copy_pos_on_last, StepState* st,
                              unsigned long long cond_handle, int use_cond, cudaStream_t stream);
// pair_ev.cu: the evaluators beyond LJ (enum B200_PAIR_* of include/hoomd_b200.h)
#define B200_DECL_EV(NAME)                                                                                              \
    cudaError_t launch_pair_##NAME(const b200_pair_args& a, const void* d_params, cudaStream_t stream);               \
    cudaError_t launch_md_step_##NAME(const b200_pair_args& pa, const void* d_params, const FusedStepArgs& fa,          \
                                      const float4* pos_in, float4* pos_out, int copy_pos_on_last, StepState* st,      \
                                      unsigned long long cond_handle, int use_cond, cudaStream_t stream);
B200_DECL_EV(gauss)
B200_DECL_EV(yukawa)
B200_DECL_EV(mie)
B200_DECL_EV(fslj)
B200_DECL_EV(morse)
B200_DECL_EV(dpd)
#undef B200_DECL_EV
// Several MD steps per launch (md_steps_persist.cuh): runs from buffer A until a rebuild is wanted or the chunk ends; the
// current positions are back in A when it returns.  cudaErrorNotSupported: not a plain LJ configuration, or the grid is not
// co-resident -- keep the one-step kernel.  probe_only: only answer that question.
cudaError_t launch_md_steps_lj(const b200_pair_args& pa, const float* d_params, const FusedStepArgs& fa, const float4* pos_a,
                               float4* pos_b, StepState* st, int probe_only, unsigned long long next_handle, int use_cond,
                               cudaStream_t stream); // use_cond: arm the conditional node `next_handle` with the exit decision
cudaError_t launch_md_step_kind(int evaluator, const b200_pair_args& pa, const void* d_params, const FusedStepArgs& fa,
                                const float4* pos_in, float4* pos_out, int copy_pos_on_last, StepState* st,
                                unsigned long long cond_handle, int use_cond, cudaStream_t stream);
// a = F/m (Integrator::computeAccelerations, hoomd/Integrator.cc:222-256)
cudaError_t launch_compute_accel(float* accel, const float4* net_force, const float4* vel, unsigned int N,
                                 cudaStream_t stream);
} // namespace b200

namespace b200
{
// bond.cu
cudaError_t launch_nlist_filter(unsigned int* n_neigh, unsigned int* nlist, const unsigned int* head_list,
                                const unsigned int* n_ex, const unsigned int* ex_list, unsigned int ex_pitch,
                                unsigned int max_ex, const unsigned int* tag, const unsigned int* rtag, unsigned int N,
                                const unsigned int* enable, cudaStream_t stream, unsigned int row_hint = 0);
cudaError_t launch_rekey_tables(unsigned int* bond_n_idx, unsigned int* bond_list_idx, const unsigned int* bond_n,
                                const unsigned int* bond_list, unsigned int bond_max, unsigned int* ex_n_idx,
                                unsigned int* ex_list_idx, const unsigned int* ex_n, const unsigned int* ex_list,
                                unsigned int ex_max, unsigned int pitch, const unsigned int* tag, const unsigned int* rtag,
                                unsigned int N, cudaStream_t stream);
cudaError_t launch_add_force(float4* net, float* net_virial, unsigned int net_pitch, const float4* f, const float* v,
                             unsigned int v_pitch, unsigned int N, bool virial, cudaStream_t stream);
cudaError_t launch_merge_xyz_w(float4* out, const float4* a, const float4* b, unsigned int N, cudaStream_t stream);
// sorted index array of the group members (ParticleGroup::getIndexArray after a particle sort): idx[k] ascending, *count
cudaError_t launch_group_index(unsigned int* idx, unsigned int* count, const unsigned char* group_flag,
                               const unsigned int* tag, unsigned int N, cudaStream_t stream);
// out[i] = by_tag[tag[i]] (a per-tag table in the current particle order)
cudaError_t launch_gather_by_tag(float* out, const float* by_tag, const unsigned int* tag, unsigned int N,
                                 cudaStream_t stream);
// host-driven Langevin step: fold the step's energy transfer *e into the tally (acc[0] += e, acc[1 + (ts & 1)] = e)
cudaError_t launch_tally_fold(double* acc, const float* e, unsigned int ts, cudaStream_t stream);
}
