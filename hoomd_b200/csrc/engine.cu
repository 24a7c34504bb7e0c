// ENGINE: host side of the device-resident system behind the b200_sys_* C ABI.
//
// The C++ classes below mirror the reference's plugin objects one to one -- ParticleData (hoomd/ParticleData.h),
// CellList (hoomd/CellList.h), NeighborListBinned (hoomd/md/NeighborList.h, NeighborListBinned.h), PotentialPairLJ
// (hoomd/md/PotentialPair.h), TwoStepNVE / TwoStepLangevin (hoomd/md/TwoStepNVE.h, TwoStepLangevin.h),
// IntegratorTwoStep (hoomd/md/IntegratorTwoStep.h), SFCPackUpdater (hoomd/SFCPackUpdater.h), ComputeThermo -- with the
// same method names and control flow (compute(ts), needsUpdating, buildHeadList, checkConditions, integrateStepOne/Two,
// prepRun, update).  Two execution modes produce the same trajectory:
//   * host-driven   : the reference's control flow, one flag read-back per checked step (NeighborListGPU.cc:93-108)
//   * device-driven : the whole step (step one + distance check, rebuild decision, conditional rebuild, pair force,
//                     step two) is a CUDA graph replayed without any host round trip; Nmax overflow (needs a host
//                     reallocation) sets a sticky flag and the chunk is replayed host-driven from a checkpoint.
// There is NO CPU fallback: every compute path launches the sm_100a kernels of this directory.
#include "common.cuh"
#include "kernels.h"
#include "comm.cuh"
#include "tile.h"
#include <string>
#include <vector>
#include <chrono>
#include <stdexcept>
#include <cmath>
#include <cstring>
#include <algorithm>
#include <cstdio>
#include <cstdlib>

namespace b200
{
#define CUDA_OK(expr)                                                                                   \
    do {                                                                                                \
        cudaError_t _e = (cudaError_t)(expr);                                                           \
        if (_e != cudaSuccess)                                                                          \
            throw std::runtime_error(std::string(#expr) + ": " + cudaGetErrorString(_e));               \
    } while (0)

template<class T> struct DevBuf
    {
    T* p = nullptr;
    size_t n = 0;
    bool owned = true;
    DevBuf() {}
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }
    void release() { if (p && owned) cudaFree(p); p = nullptr; n = 0; owned = true; }
    // non-owning view into a larger allocation (multi-GPU slab)
    void view(T* q, size_t count) { release(); p = q; n = count; owned = false; }
    void alloc(size_t count, bool zero = true)
        {
        release();
        n = count;
        CUDA_OK(cudaMalloc((void**)&p, sizeof(T) * std::max<size_t>(count, 1)));
        // cudaMemset runs on the legacy stream, which the engine's non-blocking stream does not wait for: finish it here
        if (zero)
            {
            CUDA_OK(cudaMemset(p, 0, sizeof(T) * std::max<size_t>(count, 1)));
            CUDA_OK(cudaDeviceSynchronize());
            }
        }
    // grow keeping contents (GPUArray::resize)
    void resize(size_t count)
        {
        T* q = nullptr;
        CUDA_OK(cudaMalloc((void**)&q, sizeof(T) * std::max<size_t>(count, 1)));
        CUDA_OK(cudaMemset(q, 0, sizeof(T) * std::max<size_t>(count, 1)));
        if (p) CUDA_OK(cudaMemcpy(q, p, sizeof(T) * std::min(n, count), cudaMemcpyDeviceToDevice));
        CUDA_OK(cudaDeviceSynchronize());
        if (p) cudaFree(p);
        p = q;
        n = count;
        }
    void upload(const T* h, size_t count, cudaStream_t s = 0)
        {
        CUDA_OK(cudaMemcpyAsync(p, h, sizeof(T) * count, cudaMemcpyHostToDevice, s));
        }
    void download(T* h, size_t count, cudaStream_t s = 0) const
        {
        CUDA_OK(cudaMemcpyAsync(h, p, sizeof(T) * count, cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaStreamSynchronize(s));
        }
    };

struct f3 { float x, y, z; };
struct i3 { int x, y, z; };
} // namespace b200

using namespace b200;

// a pair of CUDA events that is destroyed on every way out of the scope (exceptions included)
struct EventPair
    {
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    EventPair()
        {
        if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess)
            {
            if (e0) cudaEventDestroy(e0);
            e0 = e1 = nullptr;
            throw std::runtime_error("cudaEventCreate failed");
            }
        }
    ~EventPair()
        {
        if (e0) cudaEventDestroy(e0);
        if (e1) cudaEventDestroy(e1);
        }
    EventPair(const EventPair&) = delete;
    EventPair& operator=(const EventPair&) = delete;
    };

struct b200_system
    {
    int device = 0;
    cudaStream_t stream = 0;
    std::string error;

    // ---- ParticleData -------------------------------------------------------------------------------------------
    unsigned int N = 0, ntypes = 1;
    unsigned int Ncap = 0;   // allocated local capacity (== N on one GPU; > N under domain decomposition)
    unsigned int Gcap = 0;   // ghost capacity: ghosts live at [Ncap, Ncap + Gcap) of pos / tag / cell arrays
    b200_box box;            // local box (== global box on one GPU)
    DevBuf<float4> pos, vel, net_force, pos_alt, vel_alt, net_force_alt;
    DevBuf<f3> accel, accel_alt;
    DevBuf<i3> image, image_alt;
    DevBuf<unsigned int> tag, rtag, tag_alt;
    DevBuf<float> net_virial; // 6 x pitch
    unsigned int virial_pitch = 0;
    bool accel_set = false;
    bool compute_virial = false; // pdata flags pressure_tensor | isotropic_virial

    // ---- CellList -----------------------------------------------------------------------------------------------
    b200_dim3 dim = {1, 1, 1};
    float ghost_width[3] = {0, 0, 0};
    float nominal_width = 0.0f;
    DevBuf<unsigned int> cell_count, cell_start, cell_type, particle_cell, particle_slot, scratch_idx, cl_conditions;
    DevBuf<float4> cell_xyzf;
    bool cl_first = true, cl_force = true;
    unsigned int cl_last_computed = 0;

    // ---- NeighborList -------------------------------------------------------------------------------------------
    bool have_nlist = false;
    float r_buff = 0.0f;
    std::vector<float> h_r_cut, h_r_listsq, h_rcut_max;
    float rcut_max_max = 0.0f;
    DevBuf<float> d_r_cut, d_r_listsq, d_rcut_max, d_maxsq;
    float last_L[3] = {0.0f, 0.0f, 0.0f}; // NeighborList::m_last_L: box lengths at the last build (distance-check lambda)
    std::vector<float> h_check_table[2];  // host copies of the two rows of d_maxsq (kept alive for the async upload)
    DevBuf<unsigned int> ck_n_neigh, ck_nlist; // list checkpoint of a chunk that starts in a scaled box
    std::vector<unsigned int> h_Nmax;
    DevBuf<unsigned int> d_Nmax, nl_conditions, head_list, n_neigh, nlist, d_req_size;
    DevBuf<float4> last_pos;
    bool nl_first_compute = true;
    unsigned int nl_last_computed = 0;
    bool force_update = true;
    unsigned int every = 0, last_updated_tstep = 0, last_checked_tstep = 0;
    bool dist_check = true;              // NeighborList::m_dist_check (md.nlist.cell(dist_check=...))
    bool last_check_result = false;
    unsigned long long num_updates = 0, num_forced = 0, num_overflows = 0;

    // ---- PotentialPairLJ ----------------------------------------------------------------------------------------
    int pair_kind = 0;           // B200_PAIR_*: which EvaluatorPair* the pair force instantiates
    DevBuf<float2> d_params;     // Scalar2 evaluators (LJ, Gauss, Yukawa, force-shifted LJ)
    DevBuf<float4> d_params4;    // Scalar4 evaluators (Mie, Morse)
    DevBuf<float> d_rcutsq, d_ronsq;
    int shift_mode = 0;
    unsigned int pair_tpp = 0;

    // ---- BondData + PotentialBond + nlist exclusions (row f1) --------------------------------------------------
    // per-particle tables keyed by TAG with partner TAGS (resolved through rtag): never rebuilt on a particle sort
    unsigned int n_bond_types = 0, bond_pitch = 0, bond_max = 0;
    std::vector<unsigned int> h_bond_tags, h_bond_types; // the bond set as given (2 tags per bond)
    DevBuf<unsigned int> bond_n;          // [N] bonds per tag
    DevBuf<uint2> bond_list;              // [bond_max * pitch] (partner tag, bond type) at [k*pitch + tag]
    int bond_kind = -1;                   // -1 none, 0 FENE, 1 harmonic
    DevBuf<float4> bond_params4;
    DevBuf<float2> bond_params2;
    DevBuf<float4> bond_force, pair_force_copy;
    DevBuf<float> bond_virial, pair_virial_copy;
    DevBuf<unsigned int> bond_flags;
    bool exclusions_set = false;
    unsigned int ex_max = 0, ex_pitch = 0;
    DevBuf<unsigned int> ex_n, ex_list;   // [N], [ex_max * pitch]: partner tag at [k*pitch + tag]
    // index-keyed working copies (rows by particle index, partners as indices), refreshed after every particle sort
    bool idx_tables = true, idx_tables_dirty = true;
    DevBuf<unsigned int> bond_n_idx, ex_n_idx, ex_list_idx;
    DevBuf<uint2> bond_list_idx;

    // ---- IntegratorTwoStep + method ---------------------------------------------------------------------------
    bool have_integrator = false;
    float dt = 0.005f;
    int method = 0; // 0 none, 1 TwoStepNVE, 2 TwoStepLangevin, 3 TwoStepNVTMTK, 4 TwoStepNPTMTK
    int limit = 0, zero_force = 0;
    float limit_val = 0.0f;
    float kT = 0.0f;
    unsigned int seed_hashed = 0;
    int noiseless_t = 0;
    DevBuf<float> d_gamma;
    // TwoStepNVTMTK: {xi, eta, exp_fac, gen} lives on the device (d_nvt), checkpointed per chunk; fixed-point accumulators
    // of sum(m v^2) for the step kernel (nvt_acc, scale nvt_scale), partial sums for the stand-alone pass
    float nvt_tau = 0.0f;
    DevBuf<float> d_nvt, ck_nvt;
    DevBuf<double> nvt_partial;
    DevBuf<unsigned long long> nvt_acc;
    double nvt_scale = 1.0;
    static constexpr unsigned int NVT_BLOCKS = 592; // CTAs of the stand-alone kinetic-energy pass (4 per SM)
    // TwoStepNPTMTK (method 4): IntegratorVariables and propagator on the host, as in the reference (steps are host-driven:
    // every step changes the box and needs two thermo read-backs)
    struct Npt
        {
        float eta = 0.0f, xi = 0.0f, nuxx = 0.0f, nuyy = 0.0f, nuzz = 0.0f; // the tilt momenta stay 0 (orthorhombic boxes)
        float tau = 1.0f, tauP = 1.0f, P = 0.0f, gamma = 0.0f;
        int couple = 4;           // 0 none, 1 xy, 2 xz, 3 yz, 4 xyz (TwoStepNPTMTK::couplingMode)
        unsigned int flags = 7;   // baro_x 1 | baro_y 2 | baro_z 4
        bool nph = false, rescale_all = false;
        float mat_exp_v[6], mat_exp_r[6], mat_exp_r_int[6];
        float V = 0.0f;
        unsigned int ndof = 0;
        // the full-step ComputeThermo caches its result per time step (Compute::shouldCompute)
        bool th_valid = false;
        unsigned int th_ts = 0;
        float Pt[6] = {0, 0, 0, 0, 0, 0}, ke_t = 0.0f; // {xx, xy, xz, yy, yz, zz}
        } npt;
    DevBuf<double> npt_partial, npt_out;

    // md.pair.dpd: the DPD pair force as a thermostat (b200_sys_set_dpd_thermostat); steps are host-driven then
    bool dpd_thermo = false;
    float dpd_kT = 0.0f;
    unsigned int dpd_seed = 0;

    // ParticleGroup of the integration method (hoomd/ParticleGroup.h): 1 byte per TAG; empty = group.all().  group_idx is
    // ParticleGroup::getIndexArray (sorted member indices), rebuilt after every particle sort, for the stand-alone drivers.
    DevBuf<unsigned char> group_flag;
    DevBuf<unsigned int> group_idx, group_cnt;
    unsigned int group_size = 0;         // members (0 with an empty group_flag: all N)
    bool group_idx_dirty = true;
    // md.integrate.langevin(dscale=lambda, tally=True)
    DevBuf<float> diameter;              // per TAG (ParticleData default 1.0 when empty)
    int use_lambda = 0, tally = 0;
    float lambda = 0.0f;
    DevBuf<double> tally_acc, ck_tally;  // {running sum of E_step, E of even steps, E of odd steps}
    DevBuf<float> tally_tmp;

    // ---- SFCPackUpdater ---------------------------------------------------------------------------------------
    unsigned int sorter_period = 0, sorter_next = 0, sorter_grid = 256;
    DevBuf<unsigned int> sort_order, sort_keys, sort_scratch;
    DevBuf<unsigned char> sort_temp;

    // ---- System -------------------------------------------------------------------------------------------------
    unsigned int timestep = 0;
    int use_graph = 2;                   // 0 host-driven, 1 plain graph, 2 conditional graph (default), 3 stream launches
    unsigned int* h_flag = nullptr;      // pinned: distance-check result / small read-backs
    DevBuf<unsigned int> d_flag;
    DevBuf<StepState> d_state;
    StepState* h_state = nullptr;        // pinned
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t graph_exec = nullptr;
    bool graph_valid = false;
    int graph_uses_cond = 0;
    // MD steps per launch of the conditional graph (even; HOOMD_B200_GRAPH_UNROLL): [k_graph_pre] -> U x (IF{rebuild} -> k_md_step)
    unsigned int graph_unroll = 2, graph_steps = 2;
    // Multi-step launches (k_md_steps, md_steps_persist.cuh): every graph unit is [k_graph_pre] -> IF{rebuild} -> k_md_steps,
    // which runs until the next rebuild.  persist_mode: -1 auto (on when the configuration and the grid allow it), 0 off
    int persist_mode = -1;
    bool graph_persist = false;
    double persist_steps_per_unit = 4.0; // running estimate of the MD steps one unit advances (the rebuild interval)
    unsigned long long persist_units_last_run = 0;
    unsigned long long kernels_last_run = 0;
    // checkpoint of the device-driven chunk
    DevBuf<float4> ck_pos, ck_vel, ck_last_pos;
    DevBuf<f3> ck_accel;
    DevBuf<i3> ck_image;
    // thermo
    DevBuf<double> th_partial, th_out;

    // ---- DomainDecomposition + Communicator (multi-GPU) -------------------------------------------------------
    bool mg = false;
    b200_box global_box;
    unsigned int mg_grid[3] = {1, 1, 1}, mg_pos[3] = {0, 0, 0};
    float r_ghost = 0.0f;
    MgConst mgc;
    MgLocal mgl;
    void* slab = nullptr;            // IPC-exported: posA | posB | tag | migration mailboxes | counts | flags
    size_t slab_bytes = 0, off_posA = 0, off_posB = 0, off_tag = 0, off_mig = 0, off_migcnt = 0, off_flags = 0;
    void* peer_base[MG_MAX_RANKS] = {nullptr};
    bool mg_connected = false;
    DevBuf<MgState> d_ms;
    MgState* h_ms = nullptr;         // pinned
    DevBuf<unsigned int> holes, low_holes, hole_mark, plan_src;
    DevBuf<unsigned char> slot_dir, pplan_n;
    DevBuf<unsigned int> pplan, chunk_flag, cta_order;
    DevBuf<float> recv_shift;        // [27 * 3] periodic image shift of the ghosts received in segment d (binning only)
    DevBuf<GhostSendTable> gs_table;
    unsigned int mg_n_global = 0;    // particles of all ranks (thermostat degrees of freedom)
    unsigned int mg_tpp = 2;         // lanes per particle of k_md_step under decomposition (fixes the chunk size of cta_order)
    // Ghost update inside k_md_step with the boundary chunks scheduled first, so the NVLink stores overlap the interior force
    // work (the default).  HOOMD_B200_FUSED_GHOSTS=0: the round-1 sequence barrier kernel -> k_md_step ->
    // k_mg_ghost_send, kept for A/B runs.
    bool fused_ghosts = true;
    // HOOMD_B200_MG_MODE (A/B and debugging): bit 0 = ghost stores inside k_md_step, bit 1 = boundary-first CTA order,
    // bit 2 = the step barrier inside k_md_step's last CTA (needs bit 0).  Default 3: measured on 8 B200 (profiles/r2h_*) the
    // in-kernel barrier (7) is slower than the one-warp barrier kernel -- a rank's last CTA spins while its SMs are
    // otherwise idle either way, and the ranks reach it less aligned -- 3059 vs 3275 TPS.
    int mg_mode = 3;

    // ---- cell-tile path (tile.cu): neighbor positions from shared memory, particles kept in cell order -------------
    int pair_mode = -1;              // -1 auto (tile when supported), 0 CSR walk, 1 tile (error when unsupported)
    bool tile_on = false;
    unsigned int tile_req[3] = {0, 0, 0}; // requested cells per tile (0 = auto)
    TileGeom tile_geom = {};
    unsigned int tile_tpp = 0, tile_Nmax = 0, tile_pitch = 0;
    DevBuf<uint4> tile_ell;
    DevBuf<unsigned int> tile_nn, tile_perm, tile_cond, ck_tag;
    };

namespace b200
{
static void invalidate_graph(b200_system* s)
    {
    if (s->graph_exec) cudaGraphExecDestroy(s->graph_exec);
    if (s->graph) cudaGraphDestroy(s->graph);
    s->graph_exec = nullptr;
    s->graph = nullptr;
    s->graph_valid = false;
    }

// ---- CellList::compute (hoomd/CellList.cc:133-199) ------------------------------------------------------------------
static void cell_list_args(b200_system* s, b200_cell_list_args& a)
    {
    memset(&a, 0, sizeof(a));
    a.d_cell_start = s->cell_start.p;
    a.d_cell_xyzf = (float*)s->cell_xyzf.p;
    a.d_cell_type = s->ntypes > 1 ? s->cell_type.p : nullptr;
    a.d_cell_idx = s->tile_on ? s->tile_perm.p : nullptr;
    a.flag_type = s->tile_on ? 1 : 0; // tile path: cell_xyzf is the cell-sorted copy of pos (type in w)
    a.d_conditions = s->cl_conditions.p;
    a.d_fused_barrier = s->mg ? nullptr : s->cl_conditions.p + 4; // (ghost shifts under decomposition: four-launch build)
    a.d_cell_count = s->cell_count.p;
    a.d_particle_cell = s->particle_cell.p;
    a.d_particle_slot = s->particle_slot.p;
    a.d_scratch_idx = s->scratch_idx.p;
    a.d_pos = (const float*)s->pos.p;
    a.N = s->mg ? s->Ncap : s->N; // multi-GPU: the ghost base; the live local count is read on the device
    a.n_ghost = s->Gcap;
    a.box = s->box;
    a.dim = s->dim;
    for (int d = 0; d < 3; d++) a.ghost_width[d] = s->ghost_width[d];
    if (s->mg)
        {
        // ghosts hold raw coordinates: bin them at their periodic image next to the local box
        a.d_ghost_dir = s->slot_dir.p;
        a.d_ghost_shift = s->recv_shift.p;
        }
    }

static void nlist_args(b200_system* s, b200_nlist_args& a)
    {
    memset(&a, 0, sizeof(a));
    a.d_nlist = s->nlist.p;
    a.d_n_neigh = s->n_neigh.p;
    a.d_last_pos = (float*)s->last_pos.p;
    a.d_conditions = s->nl_conditions.p;
    a.d_Nmax = s->d_Nmax.p;
    a.d_head_list = s->head_list.p;
    a.d_pos = (const float*)s->pos.p;
    a.N = s->mg ? s->Ncap : s->N; // entries with an index >= N are ghosts
    a.d_cell_start = s->cell_start.p;
    a.d_cell_xyzf = (const float*)s->cell_xyzf.p;
    a.d_cell_type = s->ntypes > 1 ? s->cell_type.p : nullptr;
    a.dim = s->dim;
    a.box = s->box;
    for (int d = 0; d < 3; d++) a.ghost_width[d] = s->ghost_width[d];
    a.d_r_cut = s->d_r_cut.p;
    a.d_r_listsq = s->d_r_listsq.p;
    a.ntypes = s->ntypes;
    a.half_storage = 0;
    a.image_box = s->mg ? &s->global_box : nullptr; // decomposition: minimum images in the global box (raw ghost coordinates)
    // staging capacity of a build pass: the mean 27-cell population with head room for density fluctuations
    const double per_cell = (double)(s->N + (s->mg ? s->Gcap / 2 : 0)) / ((double)s->dim.x * s->dim.y * s->dim.z);
    a.cand_hint = (unsigned int)std::min(1024.0, 27.0 * per_cell * 1.5 + 48.0);
    }

static void pair_args(b200_system* s, b200_pair_args& a)
    {
    memset(&a, 0, sizeof(a));
    a.d_force = (float*)s->net_force.p; // single ForceCompute: written straight into the net force (fused K8)
    a.d_virial = s->net_virial.p;
    a.virial_pitch = s->virial_pitch;
    a.N = s->N;
    a.d_pos = (const float*)s->pos.p;
    a.box = s->mg ? s->global_box : s->box; // the box minimum images are taken in (PotentialPair.h:337: getGlobalBox())
    a.d_n_neigh = s->n_neigh.p;
    a.d_nlist = s->nlist.p;
    a.d_head_list = s->head_list.p;
    a.d_rcutsq = s->d_rcutsq.p;
    a.d_ronsq = s->d_ronsq.p;
    a.ntypes = s->ntypes;
    a.shift_mode = (unsigned int)s->shift_mode;
    a.compute_virial = s->compute_virial ? 1u : 0u;
    a.threads_per_particle = s->mg ? s->mg_tpp : s->pair_tpp; // decomposition: fixed, it shapes cta_order
    a.block_size = 0;
    // listed pairs are at most r_list apart at build time and each partner drifts < r_buff/2 (+ one step) before the
    // next rebuild: r_cut + 2 r_buff bounds every |r_i - r_j| the kernel will see
    a.wrap_skip_margin = s->rcut_max_max + 2.0f * s->r_buff + 0.1f;
    }

static const void* pair_params_ptr(b200_system* s)
    {
    return (s->pair_kind == B200_PAIR_MIE || s->pair_kind == B200_PAIR_MORSE) ? (const void*)s->d_params4.p
                                                                              : (const void*)s->d_params.p;
    }

// IntegratorTwoStep::getNDOF for one method on all particles (hoomd/md/IntegratorTwoStep.cc:270-283): D N - D
static unsigned int method_group_size(const b200_system* s) { return s->group_flag.n ? s->group_size : s->N; }
// (domain decomposition: the group is all particles of ALL ranks)
static unsigned int nvt_ndof(const b200_system* s) { return 3u * (s->mg ? s->mg_n_global : method_group_size(s)) - 3u; }

// ParticleGroup::getIndexArray of the method's group (NULL = all particles): rebuilt when the particle order changed
static const unsigned int* group_index(b200_system* s)
    {
    if (!s->group_flag.n) return nullptr;
    if (s->group_idx_dirty)
        {
        if (s->group_idx.n < s->N) s->group_idx.alloc(s->N, false);
        if (!s->group_cnt.n) s->group_cnt.alloc(1);
        CUDA_OK(launch_group_index(s->group_idx.p, s->group_cnt.p, s->group_flag.p, s->tag.p, s->N, s->stream));
        s->kernels_last_run += 1;
        s->group_idx_dirty = false;
        }
    return s->group_idx.p;
    }

// ComputeThermo (kinetic energy of the half-step velocities) + TwoStepNVTMTK::advanceThermostat, on the device
static void nvt_advance(b200_system* s)
    {
    // tagged with the timestep: the step kernel that runs step two of this update finds its factor ready
    CUDA_OK(launch_nvt_advance(s->d_nvt.p, s->nvt_partial.p, b200_system::NVT_BLOCKS, (const float*)s->vel.p,
                               group_index(s), method_group_size(s), nvt_ndof(s), s->kT, s->nvt_tau, s->dt, s->timestep,
                               s->stream));
    s->kernels_last_run += 2;
    }

static void fused_args(b200_system* s, FusedStepArgs& a)
    {
    memset(&a, 0, sizeof(a));
    a.pos = s->pos.p; a.vel = s->vel.p; a.accel = (float*)s->accel.p; a.image = (int*)s->image.p;
    a.net_force = s->net_force.p; a.last_pos = s->last_pos.p; a.tag = s->tag.p;
    a.rcut_max = s->d_rcut_max.p; a.gamma = s->d_gamma.p; a.maxsq = s->d_maxsq.p;
    a.N = s->mg ? s->Ncap : s->N; a.box = s->box; a.dt = s->dt; a.r_buff = s->r_buff;
    a.d_N = s->mg ? &s->d_state.p->N : nullptr;
    a.mg = s->mg ? ((s->fused_ghosts && s->mg_connected && (s->mg_mode & 4)) ? 2 : 1) : 0;
    a.method = s->method; a.limit = s->limit; a.limit_val = s->limit_val; a.zero_force = s->zero_force;
    a.kT = s->kT; a.seed = s->seed_hashed; a.noiseless_t = s->noiseless_t;
    for (int d = 0; d < 3; d++) a.lambda[d] = (s->mg || !s->have_nlist) ? 1.0f : s->box.L[d] / s->last_L[d];
    a.lambda_min = std::min(std::min(a.lambda[0], a.lambda[1]), a.lambda[2]);
    if (s->mg && s->fused_ghosts && s->mg_connected)
        {
        a.gs.tab = s->gs_table.p; a.gs.pplan = s->pplan.p; a.gs.pplan_n = s->pplan_n.p; a.gs.slot_dir = s->slot_dir.p;
        a.gs.pitch = s->Ncap;
        a.gs.cta_order = (s->mg_mode & 2) ? s->cta_order.p : nullptr;
        }
    if (s->bond_kind >= 0)
        {
        a.extra_force = s->bond_force.p; a.extra_virial = s->bond_virial.p; a.extra_virial_pitch = s->virial_pitch;
        }
    if (s->method == 3)
        {
        a.nvt_state = s->d_nvt.p; a.nvt_acc = s->nvt_acc.p; a.nvt_scale = s->nvt_scale; a.nvt_ndof = nvt_ndof(s);
        a.nvt_tau = s->nvt_tau;
        }
    a.group_flag = s->group_flag.n ? s->group_flag.p : nullptr;
    static const int head_arith = getenv("HOOMD_B200_HEAD_ARITH") ? atoi(getenv("HOOMD_B200_HEAD_ARITH")) : 1;
    static const int prefetch = getenv("HOOMD_B200_PREFETCH") ? atoi(getenv("HOOMD_B200_PREFETCH")) : 1;
    a.nmax_uniform = (head_arith && s->ntypes == 1 && !s->h_Nmax.empty()) ? s->h_Nmax[0] : 0u; // baked into the step graph
    a.prefetch = prefetch;
    if (s->method == 2)
        {
        a.use_lambda = s->use_lambda; a.gamma_lambda = s->lambda; a.diameter = s->diameter.p;
        a.tally_acc = s->tally ? s->tally_acc.p : nullptr;
        }
    }

// BondedGroupData::rebuildGPUTable + NeighborList::updateExListIdx: refresh the index-keyed tables when the particle order
// or the bond / exclusion sets changed (never inside a captured graph: callers run this before launching steps)
static void ensure_idx_tables(b200_system* s)
    {
    if (!s->idx_tables || !s->idx_tables_dirty) return;
    const bool have_b = s->bond_n.n != 0, have_e = s->exclusions_set && s->ex_max;
    if (have_b || have_e)
        {
        CUDA_OK(launch_rekey_tables(have_b ? s->bond_n_idx.p : nullptr, (unsigned int*)s->bond_list_idx.p,
                                    have_b ? s->bond_n.p : nullptr, (const unsigned int*)s->bond_list.p, s->bond_max,
                                    have_e ? s->ex_n_idx.p : nullptr, s->ex_list_idx.p, have_e ? s->ex_n.p : nullptr,
                                    s->ex_list.p, s->ex_max, s->N, s->tag.p, s->rtag.p, s->N, s->stream));
        s->kernels_last_run += 1;
        }
    s->idx_tables_dirty = false;
    }

// PotentialBond<evaluator>::computeForces at `pos` into bond_force / bond_virial (hoomd/md/PotentialBond.h:155-300)
static void enqueue_bond_forces(b200_system* s, const float4* pos)
    {
    if (s->bond_kind < 0) return;
    b200_bond_args a;
    memset(&a, 0, sizeof(a));
    a.d_force = (float*)s->bond_force.p; a.d_virial = s->bond_virial.p; a.virial_pitch = s->virial_pitch;
    a.N = s->N; a.d_pos = (const float*)pos; a.d_diameter = nullptr; a.box = s->box;
    if (s->mg)
        {
        // every slot of the local capacity (unused ones carry NO_TAG and are skipped: the live count changes on the device);
        // partners may be ghosts with raw coordinates: minimum images in the GLOBAL box
        a.N = s->Ncap;
        a.box = s->global_box;
        }
    a.d_bond_list = (const unsigned int*)s->bond_list.p; a.pitch = s->bond_pitch; a.d_n_bonds = s->bond_n.p;
    a.n_bond_types = s->n_bond_types; a.compute_virial = s->compute_virial ? 1u : 0u;
    a.d_tag = s->tag.p; a.d_rtag = s->rtag.p;
    if (s->idx_tables)
        {
        a.d_bond_list = (const unsigned int*)s->bond_list_idx.p; a.d_n_bonds = s->bond_n_idx.p;
        a.d_tag = nullptr; a.d_rtag = nullptr;
        }
    if (s->bond_kind == 0)
        CUDA_OK(b200_compute_fene_bond_forces(&a, (const float*)s->bond_params4.p, s->bond_flags.p, s->stream));
    else
        CUDA_OK(b200_compute_harmonic_bond_forces(&a, (const float*)s->bond_params2.p, s->bond_flags.p, s->stream));
    }

// "bond.fene: bond out of bounds" (hoomd/md/PotentialBond.h:288-292): raised at the next host synchronisation point
static void check_bond_flags(b200_system* s)
    {
    if (s->bond_kind < 0) return;
    unsigned int f = 0;
    s->bond_flags.download(&f, 1, s->stream);
    if (f)
        {
        CUDA_OK(cudaMemsetAsync(s->bond_flags.p, 0, sizeof(unsigned int), s->stream));
        if (f == 2u)
            throw std::runtime_error("bond: a bond partner is neither a local particle nor a ghost of this rank (bond longer "
                                     "than the ghost layer?)");
        throw std::runtime_error(std::string("bond.") + (s->bond_kind == 0 ? "fene" : "harmonic")
                                 + ": bond out of bounds. Error in bond calculation");
        }
    }

// NeighborList::filterNlist (hoomd/md/NeighborList.cc:246-247,1275-1328)
static void enqueue_nlist_filter(b200_system* s, const unsigned int* enable)
    {
    if (!s->exclusions_set || s->ex_max == 0) return;
    if (s->idx_tables)
        CUDA_OK(launch_nlist_filter(s->n_neigh.p, s->nlist.p, s->head_list.p, s->ex_n_idx.p, s->ex_list_idx.p, s->ex_pitch,
                                    s->ex_max, nullptr, nullptr, s->N, enable, s->stream, s->h_Nmax[0]));
    else
        CUDA_OK(launch_nlist_filter(s->n_neigh.p, s->nlist.p, s->head_list.p, s->ex_n.p, s->ex_list.p, s->ex_pitch,
                                    s->ex_max, s->tag.p, s->rtag.p, s->mg ? s->Ncap : s->N, enable, s->stream, s->h_Nmax[0]));
    }
} // namespace b200

namespace b200
{
static void tile_nlist_build(b200_system* s);
static bool tile_check_conditions(b200_system* s);
static void tile_step_args(b200_system* s, TileStepArgs& a, int force_only);
static void tile_enqueue_rebuild(b200_system* s, const unsigned int* enable, float4* pos);

// ---- CellList ---------------------------------------------------------------------------------------------------------
static void cell_list_check_conditions(b200_system* s)
    {
    unsigned int c[3];
    s->cl_conditions.download(c, 3, s->stream);
    if (c[1])
        throw std::runtime_error("Error computing cell list: particle with index " + std::to_string(c[1] - 1)
                                 + " has NaN for its position.");
    if (c[2])
        throw std::runtime_error("Error computing cell list: particle with index " + std::to_string(c[2] - 1)
                                 + " is no longer in the simulation box.");
    }

static void cell_list_compute(b200_system* s, unsigned int ts)
    {
    // Compute::shouldCompute + m_force_compute (particle sort / box change)
    bool should = s->cl_first || s->cl_force || s->cl_last_computed != ts;
    if (!should) return;
    s->cl_first = false;
    s->cl_force = false;
    s->cl_last_computed = ts;
    b200_cell_list_args a;
    cell_list_args(s, a);
    CUDA_OK(launch_cell_list(&a, nullptr, s->stream));
    s->kernels_last_run += cell_list_kernels(a.dim.x * a.dim.y * a.dim.z, a.N + a.n_ghost, a.d_fused_barrier != nullptr);
    cell_list_check_conditions(s);
    }

// ---- NeighborList -----------------------------------------------------------------------------------------------------
// Row `row` of the device distance-check table (hoomd/md/NeighborList.cc:981-1006, FP32 like the reference): per type the
// squared threshold ((r_cut_max + r_buff) lambda_min - r_cut_max) / 2, then lambda x/y/z.  Row 0: lambda = 1.
static void upload_check_table(b200_system* s, int row)
    {
    const unsigned int nt = s->ntypes;
    float lambda[3] = {1.0f, 1.0f, 1.0f};
    if (row)
        for (int d = 0; d < 3; d++) lambda[d] = s->box.L[d] / s->last_L[d];
    const float lambda_min = std::min(std::min(lambda[0], lambda[1]), lambda[2]);
    std::vector<float>& t = s->h_check_table[row];
    t.assign(nt + 3, 0.0f);
    for (unsigned int i = 0; i < nt; i++)
        {
        const float old_rmin = s->h_rcut_max[i];
        volatile float rmax = old_rmin + s->r_buff;
        volatile float num = rmax * lambda_min - old_rmin;
        const float delta_max = num / 2.0f;
        t[i] = (delta_max > 0.0f) ? delta_max * delta_max : 0.0f;
        }
    for (int d = 0; d < 3; d++) t[nt + d] = lambda[d];
    CUDA_OK(cudaMemcpyAsync(s->d_maxsq.p + (size_t)row * (nt + 3), t.data(), sizeof(float) * (nt + 3),
                            cudaMemcpyHostToDevice, s->stream));
    }

// NeighborList::updateRList (hoomd/md/NeighborList.cc:354-400)
static void nlist_update_rlist(b200_system* s)
    {
    const unsigned int nt = s->ntypes;
    s->h_r_listsq.assign(nt * nt, 0.0f);
    s->h_rcut_max.assign(nt, 0.0f);
    float r_cut_max = 0.0f;
    for (unsigned int i = 0; i < nt; i++)
        {
        float r_cut_max_i = 0.0f;
        for (unsigned int j = 0; j < nt; j++)
            {
            const float r_cut_ij = s->h_r_cut[i * nt + j];
            if (r_cut_ij > r_cut_max_i) r_cut_max_i = r_cut_ij;
            float r_list = (r_cut_ij > 0.0f) ? r_cut_ij + s->r_buff : 0.0f;
            s->h_r_listsq[i * nt + j] = r_list * r_list;
            }
        s->h_rcut_max[i] = r_cut_max_i;
        if (r_cut_max_i > r_cut_max) r_cut_max = r_cut_max_i;
        }
    s->rcut_max_max = r_cut_max;
    s->d_r_cut.upload(s->h_r_cut.data(), nt * nt, s->stream);
    s->d_r_listsq.upload(s->h_r_listsq.data(), nt * nt, s->stream);
    s->d_rcut_max.upload(s->h_rcut_max.data(), nt, s->stream);
    // distance-check table for the device-driven steps: row 0 = fixed box (lambda = 1), row 1 = box scaled since the build
    upload_check_table(s, 0);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    }

// NeighborList::buildHeadList + resizeNlist (hoomd/md/NeighborList.cc:1336-1380)
static void nlist_build_head_list(b200_system* s)
    {
    if (s->tile_on) return; // ELL rows: no head list
    // domain decomposition keeps Nmax equal for all types (nlist_check_conditions): the head list is i * Nmax whatever the
    // types of the slots are -- and the unused capacity slots hold no valid type to look up
    CUDA_OK(b200_nlist_build_head_list(s->head_list.p, s->d_req_size.p, s->d_Nmax.p, (const float*)s->pos.p,
                                       s->mg ? s->Ncap : s->N,
                                       s->mg ? 1u : s->ntypes, s->stream));
    s->kernels_last_run += (s->ntypes == 1) ? 1 : 3;
    unsigned int size = 0;
    s->d_req_size.download(&size, 1, s->stream);
    if (size > s->nlist.n)
        {
        size_t alloc_size = s->nlist.n ? s->nlist.n : 1;
        // amortised growth (the reference uses x1.125, hoomd/md/NeighborList.cc:1365-1380; a larger factor keeps the
        // captured step graph, which bakes the list's address, alive across more Nmax increases)
        while (size > alloc_size) alloc_size = ((size_t)(((float)alloc_size) * 1.25f)) + 1;
        s->nlist.alloc(alloc_size, false);
        invalidate_graph(s);
        }
    }

// NeighborList::checkConditions (hoomd/md/NeighborList.cc:1389-1405)
static bool nlist_check_conditions(b200_system* s)
    {
    std::vector<unsigned int> c(s->ntypes);
    s->nl_conditions.download(c.data(), s->ntypes, s->stream);
    bool result = false;
    for (unsigned int i = 0; i < s->ntypes; i++)
        if (c[i] > s->h_Nmax[i])
            {
            s->h_Nmax[i] = (c[i] > 8) ? (c[i] + 7) & ~7u : 8;
            result = true;
            }
    if (result)
        {
        if (s->mg)
            {
            // Domain decomposition: ONE row capacity for every type.  Local slots change their type at every migration (on the
            // device), and with equal Nmax the head list i * Nmax does not depend on the types, so it never has to be rebuilt
            // inside the rebuild bracket.  (Rows are a little longer than the reference's per-type Nmax; contents are the same.)
            const unsigned int m = *std::max_element(s->h_Nmax.begin(), s->h_Nmax.end());
            std::fill(s->h_Nmax.begin(), s->h_Nmax.end(), m);
            }
        s->d_Nmax.upload(s->h_Nmax.data(), s->ntypes, s->stream);
        CUDA_OK(cudaStreamSynchronize(s->stream));
        invalidate_graph(s); // Nmax[0] is a by-value argument of the captured step kernels (FusedStepArgs::nmax_uniform)
        }
    return result;
    }

// NeighborList::distanceCheck (hoomd/md/NeighborList.cc:964-1040): kernel + one mapped-flag style read-back
static bool nlist_distance_check(b200_system* s, unsigned int ts)
    {
    CUDA_OK(cudaMemsetAsync(s->d_flag.p, 0, sizeof(unsigned int), s->stream));
    // direction of maximum box contraction since the last build (hoomd/md/NeighborList.cc:981-986); 1 for a fixed box
    const float lambda[3] = {s->box.L[0] / s->last_L[0], s->box.L[1] / s->last_L[1], s->box.L[2] / s->last_L[2]};
    const float lambda_min = std::min(std::min(lambda[0], lambda[1]), lambda[2]);
    CUDA_OK(b200_nlist_needs_update_check(s->d_flag.p, (const float*)s->last_pos.p, (const float*)s->pos.p, s->N,
                                          &s->box, s->d_rcut_max.p, s->r_buff, s->ntypes, lambda_min, lambda, ts + 1,
                                          s->stream));
    s->kernels_last_run += 1;
    CUDA_OK(cudaMemcpyAsync(s->h_flag, s->d_flag.p, sizeof(unsigned int), cudaMemcpyDeviceToHost, s->stream));
    CUDA_OK(cudaStreamSynchronize(s->stream));
    return *s->h_flag == ts + 1;
    }

// NeighborList::needsUpdating (hoomd/md/NeighborList.cc:1080-1163)
static bool nlist_needs_updating(b200_system* s, unsigned int ts)
    {
    if (s->last_checked_tstep == ts)
        {
        if (s->force_update)
            {
            s->force_update = false;
            return true;
            }
        return s->last_check_result;
        }
    s->last_checked_tstep = ts;
    const bool should_check = !s->force_update && !(ts < s->last_updated_tstep + s->every);
    if (!s->force_update && !should_check)
        {
        s->last_check_result = false;
        return false;
        }
    bool result = false;
    if (s->force_update)
        {
        result = true;
        s->force_update = false;
        s->num_forced += 1;
        s->last_updated_tstep = ts;
        }
    else
        {
        if (s->r_buff < 1e-6f
            || (!s->dist_check && (s->every == 0 || (s->every > 1 && ts == s->last_updated_tstep + s->every))))
            result = true; // distance checks off: the check period decides (NeighborList.cc:1126-1135)
        else result = nlist_distance_check(s, ts);
        if (result)
            {
            s->last_updated_tstep = ts;
            s->num_updates += 1;
            }
        }
    s->last_check_result = result;
    return result;
    }

// NeighborListBinned::buildNlist (hoomd/md/NeighborListBinned.cc:79-240)
static void nlist_build(b200_system* s, unsigned int ts)
    {
    cell_list_compute(s, ts);
    const float rmax = s->rcut_max_max + s->r_buff;
    for (int d = 0; d < 3; d++)
        if (s->box.periodic[d] && s->box.L[d] <= rmax * 2.0f)
            throw std::runtime_error("nlist: Simulation box is too small! Particles would be interacting with themselves.");
    b200_nlist_args a;
    nlist_args(s, a);
    CUDA_OK(launch_nlist_binned(&a, nullptr, s->stream));
    s->kernels_last_run += 1;
    }

// NeighborList::compute (hoomd/md/NeighborList.cc:201-253)
static void nlist_compute(b200_system* s, unsigned int ts)
    {
    if (!s->have_nlist) throw std::runtime_error("nlist: not set up (call b200_sys_setup_lj first)");
    bool should = s->nl_first_compute || s->nl_last_computed != ts;
    s->nl_first_compute = false;
    s->nl_last_computed = ts;
    if (!should && !s->force_update) return;
    if (s->force_update) nlist_build_head_list(s);
    ensure_idx_tables(s);
    if (nlist_needs_updating(s, ts))
        {
        if (s->tile_on)
            {
            tile_nlist_build(s);
            return;
            }
        bool overflowed = false;
        do
            {
            nlist_build(s, ts);
            overflowed = nlist_check_conditions(s);
            if (overflowed)
                {
                s->num_overflows += 1;
                nlist_build_head_list(s);
                CUDA_OK(cudaMemsetAsync(s->nl_conditions.p, 0, sizeof(unsigned int) * s->ntypes, s->stream));
                }
            } while (overflowed);
        if (s->exclusions_set)
            {
            enqueue_nlist_filter(s, nullptr);
            s->kernels_last_run += 1;
            }
        // setLastUpdatedPos: done by the build kernel (like the reference GPU path, NeighborListGPUBinned.cu:280-285)
        for (int d = 0; d < 3; d++) s->last_L[d] = s->box.L[d]; // m_last_L (NeighborList.cc:1062)
        }
    }

// ---- PotentialPair::computeForces + Integrator::computeNetForce (single force => fused) -------------------------------
static void compute_net_force(b200_system* s, unsigned int ts)
    {
    nlist_compute(s, ts);
    b200_pair_args a;
    pair_args(s, a);
    if (s->tile_on)
        {
        FusedStepArgs fa;
        fused_args(s, fa);
        TileStepArgs ta;
        tile_step_args(s, ta, 1);
        CUDA_OK(launch_md_step_tile_lj(a, (const float*)s->d_params.p, fa, ta, s->pos.p, nullptr, 0, s->d_state.p, 0ull, 0,
                                       s->stream));
        }
    else if (s->dpd_thermo)
        {
        b200_dpd_thermo_args t;
        t.d_vel = (const float*)s->vel.p;
        t.d_tag = s->tag.p;
        t.seed = s->dpd_seed;
        t.timestep = ts;
        t.T = s->dpd_kT;
        t.deltaT = s->dt;
        a.threads_per_particle = 0; // the kernel's own default
        CUDA_OK(b200_compute_dpd_thermo_forces(&a, (const float*)pair_params_ptr(s), &t, s->stream));
        }
    else
        CUDA_OK(b200_compute_pair_forces(s->pair_kind, &a, pair_params_ptr(s), s->stream));
    s->kernels_last_run += 1;
    if (s->bond_kind >= 0)
        {
        // second ForceCompute + Integrator::computeNetForce: net = pair + bond.  The pair force alone is kept for
        // b200_sys_get_force_component.
        const unsigned int N = s->N;
        ensure_idx_tables(s);
        CUDA_OK(cudaMemcpyAsync(s->pair_force_copy.p, s->net_force.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
        if (s->compute_virial)
            CUDA_OK(cudaMemcpyAsync(s->pair_virial_copy.p, s->net_virial.p, sizeof(float) * 6 * (size_t)s->virial_pitch,
                                    cudaMemcpyDeviceToDevice, s->stream));
        enqueue_bond_forces(s, s->pos.p);
        CUDA_OK(launch_add_force(s->net_force.p, s->net_virial.p, s->virial_pitch, s->bond_force.p, s->bond_virial.p,
                                 s->virial_pitch, N, s->compute_virial, s->stream));
        s->kernels_last_run += 2;
        check_bond_flags(s);
        }
    }

// ---- TwoStepNPTMTK, host part (hoomd/md/TwoStepNPTMTK.cc).  Scalar is float; Variant::getValue returns double, so the
// expressions that touch m_T / m_S inline are evaluated in double exactly where the C++ promotion rules put them. ----------
static void apply_new_box(b200_system* s, const b200_box& nb);

// ComputeThermo::computeProperties with the pressure-tensor flag (hoomd/ComputeThermo.cc:224-240,300-372)
static void npt_thermo(b200_system* s, bool with_virial, float* P6, float* ke)
    {
    const unsigned int nblocks = std::max(1u, std::min(1024u, (method_group_size(s) + 255u) / 256u));
    if (s->npt_partial.n < 13u * 1024u)
        {
        s->npt_partial.alloc(13u * 1024u);
        s->npt_out.alloc(13);
        }
    CUDA_OK(b200_compute_thermo_tensor(s->npt_out.p, s->npt_partial.p, nblocks, (const float*)s->vel.p, nullptr,
                                       with_virial ? s->net_virial.p : nullptr, s->virial_pitch, group_index(s),
                                       method_group_size(s), s->stream));
    s->kernels_last_run += 2;
    double h[13];
    s->npt_out.download(h, 13, s->stream);
    const double ke_trans_total = 0.5f * (h[0] + h[3] + h[5]);
    *ke = (float)ke_trans_total;
    if (P6)
        {
        const float volume = s->box.L[0] * s->box.L[1] * s->box.L[2];
        for (int k = 0; k < 6; k++) P6[k] = (float)((h[k] + h[7 + k]) / volume);
        }
    }

// m_thermo_group_t->compute(timestep): once per time step
static void npt_thermo_full_step(b200_system* s, unsigned int ts)
    {
    b200_system::Npt& n = s->npt;
    if (n.th_valid && n.th_ts == ts) return;
    npt_thermo(s, true, n.Pt, &n.ke_t);
    n.th_valid = true;
    n.th_ts = ts;
    }

// TwoStepNPTMTK::advanceBarostat (TwoStepNPTMTK.cc:751-893)
static void npt_advance_barostat(b200_system* s, unsigned int ts)
    {
    b200_system::Npt& n = s->npt;
    npt_thermo_full_step(s, ts);
    float Pxx = n.Pt[0], Pyy = n.Pt[3], Pzz = n.Pt[5];
    bool bad = false;
    for (int k = 0; k < 6; k++) bad = bad || std::isnan(n.Pt[k]);
    const double S = (double)n.P, T = (double)s->kT;
    if (bad) Pxx = Pyy = Pzz = (float)S;
    const unsigned int d = 3;
    const float W = (float)((double)((float)(n.ndof + d) / (float)d) * T * n.tauP * n.tauP);
    float mtk_term = 2.0f * n.ke_t;
    mtk_term *= 0.5f * s->dt / (float)n.ndof / W;
    int couple = n.couple;
    if (!(n.flags & 1u))
        {
        if (couple == 4) couple = 3;
        if (couple == 1 || couple == 2) couple = 0;
        }
    if (!(n.flags & 2u))
        {
        if (couple == 4) couple = 2;
        if (couple == 3 || couple == 1) couple = 0;
        }
    if (!(n.flags & 4u))
        {
        if (couple == 4) couple = 1;
        if (couple == 3 || couple == 2) couple = 0;
        }
    float Px = Pxx, Py = Pyy, Pz = Pzz;
    if (couple == 1) Px = Py = 0.5f * (Pxx + Pyy);
    else if (couple == 2) Px = Pz = 0.5f * (Pxx + Pzz);
    else if (couple == 3) Py = Pz = 0.5f * (Pyy + Pzz);
    else if (couple == 4) Px = Py = Pz = (float)(1.0 / 3.0) * (Pxx + Pyy + Pzz);
    const float pref = 0.5f * s->dt * n.V / W;
    if (n.flags & 1u)
        {
        n.nuxx = (float)((double)n.nuxx + ((double)pref * ((double)Px - S) + (double)mtk_term));
        n.nuxx -= n.gamma * n.nuxx;
        }
    if (n.flags & 2u)
        {
        n.nuyy = (float)((double)n.nuyy + ((double)pref * ((double)Py - S) + (double)mtk_term));
        n.nuyy -= n.gamma * n.nuyy;
        }
    if (n.flags & 4u)
        {
        n.nuzz = (float)((double)n.nuzz + ((double)pref * ((double)Pz - S) + (double)mtk_term));
        n.nuzz -= n.gamma * n.nuzz;
        }
    }

// TwoStepNPTMTK::updatePropagator (TwoStepNPTMTK.cc:644-748) with nuxy = nuxz = nuyz = 0: the off-diagonal entries vanish
static void npt_update_propagator(b200_system* s)
    {
    b200_system::Npt& n = s->npt;
    static const float f_coeff[6] = {float(1.0), float(1.0 / 6.0), float(1.0 / 120.0), float(1.0 / 5040.0),
                                     float(1.0 / 362880.0), float(1.0 / 39916800.0)};
    const float dt = s->dt;
    const float nu[3] = {n.nuxx, n.nuyy, n.nuzz};
    for (int k = 0; k < 6; k++) n.mat_exp_v[k] = n.mat_exp_r[k] = n.mat_exp_r_int[k] = 0.0f;
    const int diag[3] = {0, 3, 5};
    for (int a = 0; a < 3; a++)
        {
        const float v_fac = -float(1.0 / 4.0) * nu[a];
        const float exp_v_fac_2 = expf(2.0f * v_fac * dt);
        const float r_fac = float(1.0 / 2.0) * nu[a];
        const float exp_r_fac = expf(float(1.0 / 2.0) * nu[a] * dt);
        const float exp_r_fac_2 = expf(nu[a] * dt);
        const float arg_r = r_fac * dt;
        float f_r = 0.0f, term_r = 1.0f;
        for (unsigned int i = 0; i < 6; i++)
            {
            f_r += f_coeff[i] * term_r;
            term_r = term_r * arg_r * arg_r;
            }
        n.mat_exp_v[diag[a]] = exp_v_fac_2;
        n.mat_exp_r[diag[a]] = exp_r_fac_2;
        n.mat_exp_r_int[diag[a]] = dt * exp_r_fac * f_r;
        }
    }

// TwoStepNPTMTK::advanceThermostat (TwoStepNPTMTK.cc:895-930): half-step temperature from the method's own ComputeThermo
static void npt_advance_thermostat(b200_system* s)
    {
    b200_system::Npt& n = s->npt;
    float ke;
    npt_thermo(s, false, nullptr, &ke);
    const float curr_T_trans = 2.0f / (float)n.ndof * ke;
    const float T = s->kT;
    const float xi_prime = n.xi + float(1.0 / 2.0) * s->dt / n.tau / n.tau * (curr_T_trans / T - 1.0f);
    n.xi = xi_prime + float(1.0 / 2.0) * s->dt / n.tau / n.tau * (curr_T_trans / T - 1.0f);
    n.eta += xi_prime * s->dt;
    }

// TwoStepNPTMTK::integrateStepOne (TwoStepNPTMTK.cc:177-466)
static void npt_step_one(b200_system* s, unsigned int ts)
    {
    b200_system::Npt& n = s->npt;
    if (method_group_size(s) == 0) throw std::runtime_error("integrate.npt(): Integration group empty.");
    n.V = s->box.L[0] * s->box.L[1] * s->box.L[2];
    n.ndof = nvt_ndof(s);
    npt_advance_barostat(s, ts);
    const float mtk = (n.nuxx + n.nuyy + n.nuzz) / (float)n.ndof;
    npt_update_propagator(s);
    // advance the box lengths: the lattice vectors times the (diagonal) position propagator; BoxDim::setL
    b200_box nb;
    b200_box_init(&nb, n.mat_exp_r[0] * s->box.L[0], n.mat_exp_r[3] * s->box.L[1], n.mat_exp_r[5] * s->box.L[2]);
    for (int d = 0; d < 3; d++) nb.periodic[d] = s->box.periodic[d];
    apply_new_box(s, nb);
    n.V = s->box.L[0] * s->box.L[1] * s->box.L[2];
    if (n.rescale_all)
        {
        CUDA_OK(b200_npt_mtk_rescale(s->N, (float*)s->pos.p, n.mat_exp_r[0], n.mat_exp_r[1], n.mat_exp_r[2],
                                     n.mat_exp_r[3], n.mat_exp_r[4], n.mat_exp_r[5], s->stream));
        s->kernels_last_run += 1;
        }
    const float exp_thermo_fac = expf(-float(1.0 / 2.0) * (n.xi + mtk) * s->dt);
    CUDA_OK(b200_npt_mtk_step_one((float*)s->pos.p, (float*)s->vel.p, (const float*)s->accel.p, group_index(s),
                                  method_group_size(s), exp_thermo_fac, n.mat_exp_v, n.mat_exp_r, n.mat_exp_r_int, s->dt,
                                  n.rescale_all ? 1 : 0, s->stream));
    CUDA_OK(b200_npt_mtk_wrap(s->N, (float*)s->pos.p, (int*)s->image.p, &s->box, s->stream));
    s->kernels_last_run += 2;
    if (!n.nph) npt_advance_thermostat(s);
    }

// TwoStepNPTMTK::integrateStepTwo (TwoStepNPTMTK.cc:471-580)
static void npt_step_two(b200_system* s, unsigned int ts)
    {
    b200_system::Npt& n = s->npt;
    const float mtk = (n.nuxx + n.nuyy + n.nuzz) / (float)n.ndof;
    const float exp_thermo_fac = expf(-float(1.0 / 2.0) * (n.xi + mtk) * s->dt);
    CUDA_OK(b200_npt_mtk_step_two((float*)s->vel.p, (float*)s->accel.p, group_index(s), method_group_size(s),
                                  (const float*)s->net_force.p, n.mat_exp_v, s->dt, exp_thermo_fac, s->stream));
    s->kernels_last_run += 1;
    npt_advance_barostat(s, ts + 1);
    }

static void integrate_step_one(b200_system* s, unsigned int ts)
    {
    if (s->method == 4)
        {
        npt_step_one(s, ts);
        return;
        }
    if (s->method == 3)
        {
        CUDA_OK(b200_nvt_mtk_step_one((float*)s->pos.p, (float*)s->vel.p, (const float*)s->accel.p, (int*)s->image.p,
                                      group_index(s), method_group_size(s), &s->box, s->d_nvt.p, s->dt, s->stream));
        s->kernels_last_run += 1;
        nvt_advance(s);
        return;
        }
    const bool nve = s->method == 1;
    CUDA_OK(b200_nve_step_one((float*)s->pos.p, (float*)s->vel.p, (const float*)s->accel.p, (int*)s->image.p,
                              group_index(s), method_group_size(s), &s->box, s->dt, nve ? s->limit : 0, s->limit_val,
                              nve ? s->zero_force : 0, s->stream));
    s->kernels_last_run += 1;
    }

static void integrate_step_two(b200_system* s, unsigned int ts)
    {
    if (s->method == 4)
        {
        npt_step_two(s, ts);
        return;
        }
    if (s->method == 2)
        {
        b200_langevin_args la;
        memset(&la, 0, sizeof(la));
        la.d_gamma = s->d_gamma.p;
        la.n_types = s->ntypes;
        la.T = s->kT;
        la.timestep = ts;
        la.seed = s->seed_hashed;
        la.noiseless_t = s->noiseless_t;
        la.use_lambda = s->use_lambda;
        la.lambda = s->lambda;
        la.tally = s->tally;
        la.d_sum_bdenergy = s->tally ? s->tally_tmp.p : nullptr;
        // the driver indexes diameter by particle index; the engine keeps it by tag: gather through the tag array first
        const float* diam = nullptr;
        if (s->use_lambda)
            {
            if (s->tally_tmp.n < 1 + (size_t)s->N) s->tally_tmp.resize(1 + (size_t)s->N);
            la.d_sum_bdenergy = s->tally ? s->tally_tmp.p : nullptr;
            CUDA_OK(launch_gather_by_tag(s->tally_tmp.p + 1, s->diameter.p, s->tag.p, s->N, s->stream));
            diam = s->tally_tmp.p + 1;
            s->kernels_last_run += 1;
            }
        CUDA_OK(b200_langevin_step_two((const float*)s->pos.p, (float*)s->vel.p, (float*)s->accel.p, diam, s->tag.p,
                                       group_index(s), method_group_size(s), (const float*)s->net_force.p, &la, s->dt, 3,
                                       s->stream));
        if (s->tally)
            {
            CUDA_OK(launch_tally_fold(s->tally_acc.p, s->tally_tmp.p, ts, s->stream));
            s->kernels_last_run += 1;
            }
        }
    else if (s->method == 3)
        CUDA_OK(b200_nvt_mtk_step_two((float*)s->vel.p, (float*)s->accel.p, group_index(s), method_group_size(s),
                                      (const float*)s->net_force.p, s->d_nvt.p, s->dt, s->stream));
    else
        CUDA_OK(b200_nve_step_two((float*)s->vel.p, (float*)s->accel.p, group_index(s), method_group_size(s),
                                  (const float*)s->net_force.p, s->dt, s->limit, s->limit_val, s->zero_force, s->stream));
    s->kernels_last_run += 1;
    }

// ParticleData::setGlobalBox and its box-change signal: the cell list re-dimensions itself (CellList.cc:151-171), the
// neighbor list's distance check sees lambda = L_new / L_last_build from then on
static void apply_new_box(b200_system* s, const b200_box& nb)
    {
    s->box = nb;
    s->global_box = nb;
    if (s->have_nlist)
        {
        // same bin counts => only the width changes
        b200_dim3 nd;
        b200_cell_list_dimensions(&s->box, s->nominal_width, 1, 0.0f, &nd, s->ghost_width);
        if (nd.x != s->dim.x || nd.y != s->dim.y || nd.z != s->dim.z)
            {
            CUDA_OK(cudaStreamSynchronize(s->stream));
            s->dim = nd;
            const size_t ncells = (size_t)nd.x * nd.y * nd.z;
            s->cell_count.alloc(ncells);
            s->cell_start.alloc(ncells + 1);
            }
        s->cl_force = true;
        }
    invalidate_graph(s); // the box is a by-value argument of the captured kernels
    }

// IntegratorTwoStep::prepRun (hoomd/md/IntegratorTwoStep.cc:336-398)
static void prep_run(b200_system* s, unsigned int ts)
    {
    if (!s->have_integrator || !s->method) throw std::runtime_error("integrate: no integration method set");
    compute_net_force(s, ts);
    if (!s->accel_set)
        {
        CUDA_OK(launch_compute_accel((float*)s->accel.p, s->net_force.p, s->vel.p, s->N, s->stream));
        s->kernels_last_run += 1;
        s->accel_set = true;
        }
    }

// SFCPackUpdater::update (hoomd/SFCPackUpdater.cc:70-108)
static void sorter_update(b200_system* s, bool periodic = true)
    {
    if (s->tile_on && periodic)
        {
        // the particles are re-ordered by cell at every rebuild already; keep the reference's schedule of forced list
        // rebuilds (notifyParticleSort -> NeighborList::forceUpdate) so the update counters stay comparable
        s->force_update = true;
        s->cl_force = true;
        return;
        }
    const unsigned int N = s->N, C = s->Ncap;
    if (s->sort_order.n < C)
        {
        s->sort_order.alloc(C);
        s->sort_keys.alloc(C);
        s->sort_scratch.alloc(2 * (size_t)C);
        uint64_t need = 0;
        CUDA_OK(b200_sfc_generate_sorted_order(s->sort_order.p, s->sort_keys.p, s->sort_scratch.p, nullptr, 0, &need,
                                               (const float*)s->pos.p, C, &s->box, s->sorter_grid, s->stream));
        s->sort_temp.alloc(need + 16);
        s->vel_alt.alloc(C); s->net_force_alt.alloc(C);
        s->accel_alt.alloc(C); s->image_alt.alloc(C); s->tag_alt.alloc(C);
        }
    CUDA_OK(b200_sfc_generate_sorted_order(s->sort_order.p, s->sort_keys.p, s->sort_scratch.p, s->sort_temp.p,
                                           s->sort_temp.n, nullptr, (const float*)s->pos.p, N, &s->box, s->sorter_grid,
                                           s->stream));
    b200_pdata_arrays in, out;
    memset(&in, 0, sizeof(in));
    memset(&out, 0, sizeof(out));
    in.pos = (float*)s->pos.p; in.vel = (float*)s->vel.p; in.accel = (float*)s->accel.p; in.image = (int*)s->image.p;
    in.tag = s->tag.p; in.net_force = (float*)s->net_force.p;
    out.pos = (float*)s->pos_alt.p; out.vel = (float*)s->vel_alt.p; out.accel = (float*)s->accel_alt.p;
    out.image = (int*)s->image_alt.p; out.tag = s->tag_alt.p; out.net_force = (float*)s->net_force_alt.p;
    // global tags exceed the local capacity under domain decomposition: no reverse-lookup table there
    CUDA_OK(b200_sfc_apply_sorted_order(&out, &in, s->mg ? nullptr : s->rtag.p, s->sort_order.p, N, s->stream));
    // copy back instead of swapping pointers: the captured step graph keeps its baked addresses
    CUDA_OK(cudaMemcpyAsync(s->pos.p, s->pos_alt.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
    CUDA_OK(cudaMemcpyAsync(s->vel.p, s->vel_alt.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
    CUDA_OK(cudaMemcpyAsync(s->net_force.p, s->net_force_alt.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
    CUDA_OK(cudaMemcpyAsync(s->accel.p, s->accel_alt.p, sizeof(f3) * N, cudaMemcpyDeviceToDevice, s->stream));
    CUDA_OK(cudaMemcpyAsync(s->image.p, s->image_alt.p, sizeof(i3) * N, cudaMemcpyDeviceToDevice, s->stream));
    CUDA_OK(cudaMemcpyAsync(s->tag.p, s->tag_alt.p, sizeof(unsigned int) * N, cudaMemcpyDeviceToDevice, s->stream));
    s->kernels_last_run += 4;
    // notifyParticleSort(): nlist forceUpdate, cell list force compute, bonded tables / exclusion indices rebuilt
    s->force_update = true;
    s->cl_force = true;
    s->idx_tables_dirty = true;
    s->group_idx_dirty = true;
    }
} // namespace b200

namespace b200
{
// ---- device-driven stepping -----------------------------------------------------------------------------------------------
// A chunk of n IntegratorTwoStep::update() calls (hoomd/md/IntegratorTwoStep.cc:96-161) runs as
//   prologue : k_fused_step_one(ts) + k_decide                                   (2 launches per chunk)
//   graph    : [k_graph_pre] -> IF{rebuild from A} -> k_md_step(A->B) -> IF{rebuild from B} -> k_md_step(B->A)
// replayed ceil(n/2) times.  k_md_step = force(ts+1) + step two(ts) + step one(ts+1) + distance check + rebuild decision
// in ONE launch (see fused_step.cu); positions ping-pong between pos (A) and pos_alt (B) and always end in A.
// mode 2 brackets the rebuild in conditional IF nodes (CUDA >= 12.4); mode 1 launches the rebuild kernels every step and
// lets them return at once on !need_rebuild (the `enable` pointer).
static void enqueue_rebuild(b200_system* s, const unsigned int* enable, const float4* pos)
    {
    if (s->tile_on)
        {
        tile_enqueue_rebuild(s, enable, const_cast<float4*>(pos));
        CUDA_OK(launch_tile_post(s->d_state.p, s->tile_cond.p, s->tile_Nmax, s->stream));
        return;
        }
    b200_cell_list_args ca;
    cell_list_args(s, ca);
    ca.d_pos = (const float*)pos;
    CUDA_OK(launch_cell_list(&ca, enable, s->stream));
    b200_nlist_args na;
    nlist_args(s, na);
    na.d_pos = (const float*)pos;
    // the build kernel raises StepState::overflow itself and stops the chunk (OverflowHook): no separate check kernel
    OverflowHook hook;
    hook.overflow = &s->d_state.p->overflow;
    hook.steps_left = &s->d_state.p->steps_left;
    CUDA_OK(launch_nlist_binned(&na, enable, s->stream, hook));
    // filterNlist runs on the complete rows; after an overflow the rows are discarded anyway (checkpoint replay)
    enqueue_nlist_filter(s, enable);
    }

static void enqueue_md_step(b200_system* s, const FusedStepArgs& fa, int phase, unsigned long long handle, int use_cond)
    {
    b200_pair_args pa;
    pair_args(s, pa);
    const float4* pin = phase == 0 ? s->pos.p : s->pos_alt.p;
    float4* pout = phase == 0 ? s->pos_alt.p : s->pos.p;
    enqueue_bond_forces(s, pin); // second ForceCompute at the positions this step reads; k_md_step adds it (fa.extra_force)
    if (s->tile_on)
        {
        TileStepArgs ta;
        tile_step_args(s, ta, 0);
        CUDA_OK(launch_md_step_tile_lj(pa, (const float*)s->d_params.p, fa, ta, pin, pout, phase == 1, s->d_state.p, handle,
                                       use_cond, s->stream));
        return;
        }
    CUDA_OK(launch_md_step_kind(s->pair_kind, pa, pair_params_ptr(s), fa, pin, pout, phase == 1, s->d_state.p, handle,
                                use_cond, s->stream));
    }

#if (CUDART_VERSION >= 12040)
// splice a conditional IF node in after what has been captured so far on the stream; returns its body graph
static cudaGraph_t add_if_node(b200_system* s, cudaGraph_t graph, cudaGraphConditionalHandle handle)
    {
    cudaStreamCaptureStatus status;
    const cudaGraphNode_t* deps = nullptr;
    size_t ndeps = 0;
    cudaGraph_t cap_graph = nullptr;
    CUDA_OK(cudaStreamGetCaptureInfo_v2(s->stream, &status, nullptr, &cap_graph, &deps, &ndeps));
    cudaGraphNodeParams np = {cudaGraphNodeTypeConditional};
    np.type = cudaGraphNodeTypeConditional;
    np.conditional.handle = handle;
    np.conditional.type = cudaGraphCondTypeIf;
    np.conditional.size = 1;
    cudaGraphNode_t cond_node;
    CUDA_OK(cudaGraphAddNode(&cond_node, graph, deps, ndeps, &np));
    CUDA_OK(cudaStreamUpdateCaptureDependencies(s->stream, &cond_node, 1, cudaStreamSetCaptureDependencies));
    return np.conditional.phGraph_out[0];
    }
#endif

static bool build_graph_conditional(b200_system* s)
    {
#if (CUDART_VERSION >= 12040)
    FusedStepArgs fa;
    fused_args(s, fa);
    const unsigned int* enable = &s->d_state.p->need_rebuild;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    bool capturing = false;
    try
        {
        CUDA_OK(cudaGraphCreate(&graph, 0));
        // multi-step launches: plain LJ configurations whose whole grid is co-resident
        bool persist = false;
        if (s->persist_mode != 0 && s->pair_kind == B200_PAIR_LJ && !s->tile_on && s->bond_kind < 0)
            {
            b200_pair_args ppa;
            pair_args(s, ppa);
            persist = launch_md_steps_lj(ppa, (const float*)s->d_params.p, fa, s->pos.p, s->pos_alt.p, s->d_state.p, 1, 0ull, 0,
                                         s->stream) == cudaSuccess;
            cudaGetLastError();
            }
        s->graph_persist = persist;
        const unsigned int U = persist ? 8u : std::max(2u, std::min(32u, s->graph_unroll & ~1u));
        std::vector<cudaGraphConditionalHandle> h(U);
        std::vector<cudaGraph_t> body(U);
        for (unsigned int k = 0; k < U; k++)
            CUDA_OK(cudaGraphConditionalHandleCreate(&h[k], graph, 0, cudaGraphCondAssignDefault));
        CUDA_OK(cudaStreamBeginCaptureToGraph(s->stream, graph, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
        capturing = true;
        if (!persist) CUDA_OK(launch_graph_pre(s->d_state.p, (unsigned long long)h[0], s->stream));
        for (unsigned int k = 0; k < U; k++)
            {
            if (persist)
                {
                // unit = rebuild from buffer A if wanted, then step until the next rebuild is wanted (the kernel leaves the
                // current positions in A and arms the next unit's IF node; the first one is armed from the pending decision)
                if (k == 0) CUDA_OK(launch_graph_pre(s->d_state.p, (unsigned long long)h[0], s->stream));
                body[k] = add_if_node(s, graph, h[k]);
                b200_pair_args ppa;
                pair_args(s, ppa);
                CUDA_OK(launch_md_steps_lj(ppa, (const float*)s->d_params.p, fa, s->pos.p, s->pos_alt.p, s->d_state.p, 0,
                                           k + 1 < U ? (unsigned long long)h[k + 1] : 0ull, k + 1 < U ? 1 : 0, s->stream));
                continue;
                }
            // every step kernel arms the IF node of the step behind it; the last one leaves that to the next launch's pre
            body[k] = add_if_node(s, graph, h[k]);
            if (k + 1 < U)
                enqueue_md_step(s, fa, (int)(k & 1u), (unsigned long long)h[k + 1], 1);
            else
                enqueue_md_step(s, fa, (int)(k & 1u), 0ull, 0);
            }
        cudaGraph_t out = nullptr;
        CUDA_OK(cudaStreamEndCapture(s->stream, &out));
        capturing = false;
        // bodies of the IF nodes
        for (unsigned int k = 0; k < U; k++)
            {
            CUDA_OK(cudaStreamBeginCaptureToGraph(s->stream, body[k], nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
            capturing = true;
            enqueue_rebuild(s, enable, (!persist && (k & 1u)) ? s->pos_alt.p : s->pos.p);
            CUDA_OK(cudaStreamEndCapture(s->stream, &out));
            capturing = false;
            }
        CUDA_OK(cudaGraphInstantiate(&exec, graph, 0));
        s->graph_steps = U;
        }
    catch (const std::exception& e)
        {
        if (capturing)
            {
            cudaGraph_t junk = nullptr;
            cudaStreamEndCapture(s->stream, &junk);
            }
        if (graph) cudaGraphDestroy(graph);
        cudaGetLastError();
        s->error = std::string("conditional graph unavailable, using early-exit kernels: ") + e.what();
        return false;
        }
    s->graph = graph;
    s->graph_exec = exec;
    s->graph_uses_cond = 1;
    s->graph_valid = true;
    return true;
#else
    return false;
#endif
    }

static void build_graph_plain(b200_system* s)
    {
    FusedStepArgs fa;
    fused_args(s, fa);
    const unsigned int* enable = &s->d_state.p->need_rebuild;
    CUDA_OK(cudaStreamBeginCapture(s->stream, cudaStreamCaptureModeThreadLocal));
    try
        {
        enqueue_rebuild(s, enable, s->pos.p);
        enqueue_md_step(s, fa, 0, 0ull, 0);
        enqueue_rebuild(s, enable, s->pos_alt.p);
        enqueue_md_step(s, fa, 1, 0ull, 0);
        }
    catch (...)
        {
        cudaGraph_t junk = nullptr;
        cudaStreamEndCapture(s->stream, &junk);
        if (junk) cudaGraphDestroy(junk);
        throw;
        }
    CUDA_OK(cudaStreamEndCapture(s->stream, &s->graph));
    CUDA_OK(cudaGraphInstantiate(&s->graph_exec, s->graph, 0));
    s->graph_steps = 2;
    s->graph_persist = false;
    s->graph_uses_cond = 0;
    s->graph_valid = true;
    }

static void ensure_graph(b200_system* s)
    {
    if (s->graph_valid) return;
    invalidate_graph(s);
    if (s->use_graph >= 2 && build_graph_conditional(s)) return;
    build_graph_plain(s);
    }

static bool box_scaled_since_build(const b200_system* s)
    {
    return s->have_nlist && (s->box.L[0] != s->last_L[0] || s->box.L[1] != s->last_L[1] || s->box.L[2] != s->last_L[2]);
    }

static void host_step(b200_system* s)
    {
    integrate_step_one(s, s->timestep);
    compute_net_force(s, s->timestep + 1);
    integrate_step_two(s, s->timestep);
    s->timestep++;
    }

static void run_chunk_device(b200_system* s, unsigned int chunk)
    {
    const unsigned int N = s->N;
    if (s->force_update) nlist_build_head_list(s); // types may have been permuted by the sorter
    ensure_idx_tables(s);
    // a short chunk behind an invalidated graph (update.box_resize with a small period re-sets the by-value box argument
    // every few steps): launching its dozen kernels directly is cheaper than capturing and instantiating a graph for it
    const bool direct = s->use_graph == 3 || (!s->graph_valid && chunk <= 16);
    if (!direct) ensure_graph(s);
    if (s->ck_pos.n < N)
        {
        s->ck_pos.alloc(N); s->ck_vel.alloc(N); s->ck_last_pos.alloc(N); s->ck_accel.alloc(N); s->ck_image.alloc(N);
        }
    if (s->tile_on)
        {
        // rebuilds re-order the particles: the checkpoint carries the tags too
        if (s->ck_tag.n < N) s->ck_tag.alloc(N);
        CUDA_OK(cudaMemcpyAsync(s->ck_tag.p, s->tag.p, sizeof(unsigned int) * N, cudaMemcpyDeviceToDevice, s->stream));
        }
    // checkpoint (device to device, ~72 B/particle per chunk)
    CUDA_OK(cudaMemcpyAsync(s->ck_pos.p, s->pos.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
    CUDA_OK(cudaMemcpyAsync(s->ck_vel.p, s->vel.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
    CUDA_OK(cudaMemcpyAsync(s->ck_last_pos.p, s->last_pos.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
    CUDA_OK(cudaMemcpyAsync(s->ck_accel.p, s->accel.p, sizeof(f3) * N, cudaMemcpyDeviceToDevice, s->stream));
    CUDA_OK(cudaMemcpyAsync(s->ck_image.p, s->image.p, sizeof(i3) * N, cudaMemcpyDeviceToDevice, s->stream));
    const bool ck_force_update = s->force_update;
    if (s->method == 3)
        {
        CUDA_OK(cudaMemcpyAsync(s->ck_nvt.p, s->d_nvt.p, sizeof(float) * 4, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_OK(cudaMemsetAsync(s->nvt_acc.p, 0, sizeof(unsigned long long) * s->nvt_acc.n, s->stream));
        }
    if (s->method == 2 && s->tally)
        {
        // the first k_md_step of the chunk runs step two of update(timestep): its E slot starts at zero
        CUDA_OK(cudaMemsetAsync(s->tally_acc.p + 1u + (s->timestep & 1u), 0, sizeof(double), s->stream));
        CUDA_OK(cudaMemcpyAsync(s->ck_tally.p, s->tally_acc.p, sizeof(double) * 3, cudaMemcpyDeviceToDevice, s->stream));
        }

    // The box changed since the last list build (update.box_resize): the distance checks of this chunk use the lambda-scaled
    // row of the table until the first rebuild decision clears StepState::box_scaled.  The list in force was built in the OLD
    // box, so it cannot be re-created from last_pos after an Nmax overflow: checkpoint the list itself for that case.
    const bool ck_scaled = box_scaled_since_build(s);
    if (ck_scaled)
        {
        upload_check_table(s, 1);
        if (s->ck_n_neigh.n < N) s->ck_n_neigh.alloc(N, false);
        if (s->ck_nlist.n < s->nlist.n) s->ck_nlist.alloc(s->nlist.n, false);
        CUDA_OK(cudaMemcpyAsync(s->ck_n_neigh.p, s->n_neigh.p, sizeof(unsigned int) * N, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_OK(cudaMemcpyAsync(s->ck_nlist.p, s->nlist.p, sizeof(unsigned int) * s->nlist.n, cudaMemcpyDeviceToDevice,
                                s->stream));
        }

    StepState& st = *s->h_state;
    memset(&st, 0, sizeof(st));
    st.timestep = s->timestep;
    st.force_update = s->force_update ? 1u : 0u;
    st.last_updated = s->last_updated_tstep;
    st.every = s->every;
    st.r_buff_tiny = (s->r_buff < 1e-6f) ? 1u : 0u;
    st.ntypes = s->ntypes;
    st.steps_left = chunk;
    st.box_scaled = ck_scaled ? 1u : 0u;
    st.no_dist_check = s->dist_check ? 0u : 1u;
    CUDA_OK(cudaMemcpyAsync(s->d_state.p, s->h_state, sizeof(StepState), cudaMemcpyHostToDevice, s->stream));
    FusedStepArgs fa;
    fused_args(s, fa);
    CUDA_OK(launch_fused_step_one(fa, s->d_state.p, s->stream));
    if (s->method == 3) nvt_advance(s); // thermostat of the prologue's step one; the later ones run inside k_md_step
    CUDA_OK(launch_decide(s->d_state.p, s->stream));
    // kernels behind the end of the chunk return at once (steps_left == 0): a chunk need not be a multiple of the graph
    const unsigned int per_launch = direct ? 2u : s->graph_steps;
    const unsigned int nlaunch = (chunk + per_launch - 1u) / per_launch;
    const unsigned int nreplay = nlaunch * (per_launch / 2u); // pairs of step kernels launched
    if (direct)
        {
        // same kernel sequence as the plain graph, launched directly on the stream (profilers see every kernel)
        const unsigned int* enable = &s->d_state.p->need_rebuild;
        for (unsigned int k = 0; k < nreplay; k++)
            {
            enqueue_rebuild(s, enable, s->pos.p);
            enqueue_md_step(s, fa, 0, 0ull, 0);
            enqueue_rebuild(s, enable, s->pos_alt.p);
            enqueue_md_step(s, fa, 1, 0ull, 0);
            }
        }
    else if (s->graph_persist)
        {
        // every unit advances a data-dependent number of steps (until the next rebuild): launch what the running estimate
        // of the rebuild interval asks for plus head room (an idle unit costs three empty launches), look at steps_left, and
        // top up in the rare case it was not enough
        unsigned int todo = chunk;
        double per_unit = std::max(1.0, 0.8 * s->persist_steps_per_unit);
        s->persist_units_last_run = 0;
        while (true)
            {
            const unsigned int units = (unsigned int)std::ceil(todo / per_unit) + 2u;
            const unsigned int nl = (units + s->graph_steps - 1u) / s->graph_steps;
            for (unsigned int k = 0; k < nl; k++) CUDA_OK(cudaGraphLaunch(s->graph_exec, s->stream));
            s->persist_units_last_run += (unsigned long long)nl * s->graph_steps;
            CUDA_OK(cudaMemcpyAsync(s->h_state, s->d_state.p, sizeof(StepState), cudaMemcpyDeviceToHost, s->stream));
            CUDA_OK(cudaStreamSynchronize(s->stream));
            if (st.steps_left == 0u || st.overflow || st.abort || st.nvt_error) break;
            todo = st.steps_left;
            per_unit = std::max(1.0, 0.5 * per_unit);
            }
        if (!st.overflow && !st.nvt_error)
            {
            const double seen = (double)chunk / (double)(st.num_updates + st.num_forced + 1u);
            s->persist_steps_per_unit = 0.5 * s->persist_steps_per_unit + 0.5 * seen;
            }
        }
    else
        for (unsigned int k = 0; k < nlaunch; k++) CUDA_OK(cudaGraphLaunch(s->graph_exec, s->stream));
    CUDA_OK(cudaMemcpyAsync(s->h_state, s->d_state.p, sizeof(StepState), cudaMemcpyDeviceToHost, s->stream));
    CUDA_OK(cudaStreamSynchronize(s->stream));
    if (st.nvt_error == 3u)
        {
        // the ticket line holds the arrivals of the step that never completed: clear it and the release words (epoch 0 again)
        CUDA_OK(cudaMemsetAsync(reinterpret_cast<char*>(s->d_state.p) + STEP_SYNC_OFFSET, 0,
                                (size_t)(1u + STEP_SYNC_COPIES) * STEP_SYNC_LINE, s->stream));
        CUDA_OK(cudaStreamSynchronize(s->stream));
        throw std::runtime_error("integrate: a CTA of the multi-step kernel never reached the grid barrier");
        }
    if (st.nvt_error)
        throw std::runtime_error(st.nvt_error == 1u
                                     ? "integrate.nvt: kinetic energy is more than 2^20 times the set point (diverged?)"
                                     : "integrate.nvt: thermostat hand-off inside the step kernel timed out");
    if (!st.overflow) cell_list_check_conditions(s);

    if (st.overflow)
        {
        CUDA_OK(cudaMemsetAsync(s->cl_conditions.p, 0, sizeof(unsigned int) * 3, s->stream));
        // Nmax overflow needs a host reallocation (NeighborList.cc:226-244): grow Nmax from the counts the overflowing
        // rebuild recorded -- the very values checkConditions would have seen --, restore the checkpoint and run the chunk
        // again.  The trajectory is deterministic, so the replay is exact.
        s->num_overflows += 1;
        if (ck_scaled)
            {
            // put the checkpoint back, list included (row layout and Nmax untouched), and run the chunk with the reference's
            // host protocol, which grows Nmax at the overflowing build itself
            CUDA_OK(cudaMemcpyAsync(s->pos.p, s->ck_pos.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
            CUDA_OK(cudaMemcpyAsync(s->vel.p, s->ck_vel.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
            CUDA_OK(cudaMemcpyAsync(s->last_pos.p, s->ck_last_pos.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
            CUDA_OK(cudaMemcpyAsync(s->accel.p, s->ck_accel.p, sizeof(f3) * N, cudaMemcpyDeviceToDevice, s->stream));
            CUDA_OK(cudaMemcpyAsync(s->image.p, s->ck_image.p, sizeof(i3) * N, cudaMemcpyDeviceToDevice, s->stream));
            CUDA_OK(cudaMemcpyAsync(s->n_neigh.p, s->ck_n_neigh.p, sizeof(unsigned int) * N, cudaMemcpyDeviceToDevice, s->stream));
            CUDA_OK(cudaMemcpyAsync(s->nlist.p, s->ck_nlist.p, sizeof(unsigned int) * s->nlist.n, cudaMemcpyDeviceToDevice,
                                    s->stream));
            if (s->method == 3)
                CUDA_OK(cudaMemcpyAsync(s->d_nvt.p, s->ck_nvt.p, sizeof(float) * 4, cudaMemcpyDeviceToDevice, s->stream));
            if (s->method == 2 && s->tally)
                CUDA_OK(cudaMemcpyAsync(s->tally_acc.p, s->ck_tally.p, sizeof(double) * 3, cudaMemcpyDeviceToDevice, s->stream));
            CUDA_OK(cudaMemsetAsync(s->nl_conditions.p, 0, sizeof(unsigned int) * s->ntypes, s->stream));
            s->force_update = ck_force_update;
            s->cl_force = true;
            for (unsigned int k = 0; k < chunk; k++) host_step(s);
            return;
            }
        const bool grew = !s->tile_on && nlist_check_conditions(s);
        if (grew) nlist_build_head_list(s);
        CUDA_OK(cudaMemcpyAsync(s->vel.p, s->ck_vel.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_OK(cudaMemcpyAsync(s->accel.p, s->ck_accel.p, sizeof(f3) * N, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_OK(cudaMemcpyAsync(s->image.p, s->ck_image.p, sizeof(i3) * N, cudaMemcpyDeviceToDevice, s->stream));
        if (s->method == 3)
            CUDA_OK(cudaMemcpyAsync(s->d_nvt.p, s->ck_nvt.p, sizeof(float) * 4, cudaMemcpyDeviceToDevice, s->stream));
        if (s->method == 2 && s->tally)
            CUDA_OK(cudaMemcpyAsync(s->tally_acc.p, s->ck_tally.p, sizeof(double) * 3, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_OK(cudaMemsetAsync(s->nl_conditions.p, 0, sizeof(unsigned int) * s->ntypes, s->stream));
        s->force_update = ck_force_update;
        if (s->tile_on)
            {
            CUDA_OK(cudaMemcpyAsync(s->tag.p, s->ck_tag.p, sizeof(unsigned int) * N, cudaMemcpyDeviceToDevice, s->stream));
            // grow the capacity that overflowed (the device rebuild left its high-water marks in tile_cond)
            tile_check_conditions(s);
            if (!ck_force_update)
                {
                // the checkpointed order IS the cell order of last_pos: this rebuild is the identity permutation
                CUDA_OK(launch_merge_xyz_w(s->pos.p, s->ck_last_pos.p, s->ck_pos.p, N, s->stream));
                tile_nlist_build(s);
                }
            }
        else if (!ck_force_update)
            {
            // re-create the list that was current at the checkpoint: it is a pure function of last_pos
            CUDA_OK(launch_merge_xyz_w(s->pos.p, s->ck_last_pos.p, s->ck_pos.p, N, s->stream));
            bool overflowed;
            do
                {
                s->cl_force = true;
                nlist_build(s, s->timestep);
                overflowed = nlist_check_conditions(s);
                if (overflowed)
                    {
                    nlist_build_head_list(s);
                    CUDA_OK(cudaMemsetAsync(s->nl_conditions.p, 0, sizeof(unsigned int) * s->ntypes, s->stream));
                    }
                } while (overflowed);
            enqueue_nlist_filter(s, nullptr);
            }
        CUDA_OK(cudaMemcpyAsync(s->pos.p, s->ck_pos.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
        s->cl_force = true;
        if (grew)
            run_chunk_device(s, chunk); // device-driven again with the larger rows (recurses on a further overflow)
        else
            for (unsigned int k = 0; k < chunk; k++) host_step(s); // the reference's host protocol, step by step
        return;
        }
    check_bond_flags(s);
    const unsigned long long rebuilds = (unsigned long long)st.num_updates + st.num_forced;
    if (rebuilds)
        for (int d = 0; d < 3; d++) s->last_L[d] = s->box.L[d]; // setLastUpdatedPos of the device-side builds
    const long long cl_extra = (long long)cell_list_kernels(s->dim.x * s->dim.y * s->dim.z, s->N, true) - 4;
    const unsigned long long rk = (unsigned long long)((s->tile_on ? 8ll : 5ll) + cl_extra)
                                  + ((s->exclusions_set && s->ex_max) ? 1ull : 0ull); // kernels per rebuild
    if (s->bond_kind >= 0) s->kernels_last_run += 2ull * nreplay;
    if (s->graph_persist && !direct)
        s->kernels_last_run += 2ull + s->persist_units_last_run + s->persist_units_last_run / s->graph_steps
                               + rk * rebuilds; // k_md_steps per unit, k_graph_pre per graph launch
    else
        s->kernels_last_run += 2ull + ((s->graph_uses_cond && !direct) ? (2ull * nreplay + nlaunch + rk * rebuilds)
                                                                                 : (2ull + 2ull * rk) * nreplay);
    s->timestep = st.timestep;
    s->last_updated_tstep = st.last_updated;
    s->num_updates += st.num_updates;
    s->num_forced += st.num_forced;
    s->force_update = st.force_update != 0;
    s->last_checked_tstep = st.timestep;
    s->last_check_result = st.need_rebuild != 0;
    s->nl_last_computed = st.timestep;
    s->nl_first_compute = false;
    if (rebuilds)
        {
        s->cl_last_computed = st.last_updated;
        s->cl_first = false;
        s->cl_force = false;
        }
    }
} // namespace b200

#include "engine_mg.inl"
#include "engine_tile.inl"

namespace b200
{
static void sys_open_device(b200_system* s)
    {
    int ndev = 0;
    CUDA_OK(cudaGetDeviceCount(&ndev));
    if (ndev == 0) throw std::runtime_error("no CUDA device: hoomd_b200 has no CPU fallback");
    CUDA_OK(cudaSetDevice(s->device));
    CUDA_OK(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    // developer overrides of the pair path: HOOMD_B200_PAIR_MODE = -1 | 0 | 1, HOOMD_B200_TILE = "bx,by,bz,tpp"
    if (const char* m = getenv("HOOMD_B200_PAIR_MODE")) s->pair_mode = atoi(m);
    if (const char* m = getenv("HOOMD_B200_FUSED_GHOSTS")) s->fused_ghosts = atoi(m) != 0;
    if (const char* m = getenv("HOOMD_B200_MG_MODE")) s->mg_mode = atoi(m);
    if (!s->fused_ghosts) s->mg_mode = 0;
    s->fused_ghosts = (s->mg_mode & 1) != 0;
    if (!s->fused_ghosts) s->mg_mode = 0;
    if (const char* m = getenv("HOOMD_B200_TAG_TABLES")) s->idx_tables = atoi(m) == 0; // A/B: tag-keyed bond / exclusion rows
    if (const char* t = getenv("HOOMD_B200_TILE"))
        {
        unsigned int v[4] = {0, 0, 0, 0};
        sscanf(t, "%u,%u,%u,%u", &v[0], &v[1], &v[2], &v[3]);
        s->tile_req[0] = v[0]; s->tile_req[1] = v[1]; s->tile_req[2] = v[2];
        s->tile_tpp = v[3];
        }
    }

// per-particle arrays sized for the local capacity; pos / pos_alt / tag are allocated by the caller (they carry the
// ghost region under domain decomposition)
static void sys_alloc_common(b200_system* s)
    {
    // Small systems are bound by the gaps between graph launches (6.9 us per step at 64k with two steps per launch): 16 steps
    // per launch, 28.8k -> 30.7k TPS (profiles/r2j_*).  Large systems gain nothing and keep the short graph.
    s->graph_unroll = s->N <= 150000u ? 16u : 2u;
    if (const char* e = getenv("HOOMD_B200_PERSIST")) s->persist_mode = atoi(e) ? -1 : 0;
    if (const char* e = getenv("HOOMD_B200_GRAPH_UNROLL")) s->graph_unroll = (unsigned int)std::max(2, atoi(e));
    const unsigned int N = s->N, C = s->Ncap;
    s->vel.alloc(C); s->net_force.alloc(C); s->accel.alloc(C); s->image.alloc(C);
    s->rtag.alloc(C); s->last_pos.alloc(C);
    s->virial_pitch = C + (16 - (C & 15)); // GPUArray 2D pitch rule (hoomd/GPUArray.h:424-425)
    s->net_virial.alloc(6 * (size_t)s->virial_pitch);
    std::vector<unsigned int> ident(C);
    for (unsigned int i = 0; i < C; i++) ident[i] = i;
    s->tag.upload(ident.data(), N, s->stream);
    s->rtag.upload(ident.data(), N, s->stream);
    // default mass 1 (ParticleData ctor)
    std::vector<float4> v(C, make_float4(0.0f, 0.0f, 0.0f, 1.0f));
    s->vel.upload(v.data(), C, s->stream);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    s->d_flag.alloc(4);
    s->d_state.alloc(STEP_STATE_ALLOC); // slot 0 + the barrier lines of k_md_steps
    s->d_req_size.alloc(1);
    s->cl_conditions.alloc(8); // {max occupancy, NaN, out of box} + pad + the two barrier words of the one-launch build
    CUDA_OK(cudaHostAlloc((void**)&s->h_flag, 4 * sizeof(unsigned int), cudaHostAllocDefault));
    CUDA_OK(cudaHostAlloc((void**)&s->h_state, sizeof(StepState), cudaHostAllocDefault));
    s->th_partial.alloc(9 * 1024);
    s->th_out.alloc(16);
    }
} // namespace b200

// ---- C ABI --------------------------------------------------------------------------------------------------------------
#define SYS_TRY(s) (s)->error.clear(); try { CUDA_OK(cudaSetDevice((s)->device));
#define SYS_CATCH(s) } catch (const std::exception& e) { (s)->error = e.what(); cudaGetLastError(); return 1; } return 0;

extern "C" {

b200_system* b200_sys_create(unsigned int N, const b200_box* box, unsigned int ntypes, int device)
    {
    b200_system* s = new b200_system;
    s->device = device;
    try
        {
        sys_open_device(s);
        s->N = N;
        s->Ncap = N;
        s->Gcap = 0;
        s->ntypes = ntypes ? ntypes : 1;
        s->box = *box;
        s->global_box = *box;
        s->pos.alloc(N);
        s->pos_alt.alloc(N); // ping-pong partner of pos (device-driven stepping) and the sorter's gather target
        s->tag.alloc(N);
        sys_alloc_common(s);
        }
    catch (const std::exception& e)
        {
        s->error = e.what();
        }
    return s;
    }

b200_system* b200_sys_create_mg(const b200_mg_params* p)
    {
    b200_system* s = new b200_system;
    s->device = p->device;
    try
        {
        sys_open_device(s);
        s->ntypes = p->ntypes ? p->ntypes : 1;
        mg_setup(s, *p);
        sys_alloc_common(s);
        mg_bind_local(s);
        }
    catch (const std::exception& e)
        {
        s->error = e.what();
        }
    return s;
    }

int b200_sys_mg_export(b200_system* s, unsigned char* handle64, b200_box* local_box, unsigned int* caps3)
    {
    SYS_TRY(s)
    if (!s->mg) throw std::runtime_error("comm: not a domain-decomposed system");
    cudaIpcMemHandle_t h;
    CUDA_OK(cudaIpcGetMemHandle(&h, s->slab));
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    memcpy(handle64, &h, 64);
    if (local_box) *local_box = s->box;
    if (caps3)
        {
        caps3[0] = s->Ncap;
        caps3[1] = s->Gcap;
        caps3[2] = s->mgc.mig_cap;
        }
    SYS_CATCH(s)
    }

int b200_sys_mg_connect(b200_system* s, const unsigned char* handles)
    {
    SYS_TRY(s)
    if (!s->mg) throw std::runtime_error("comm: not a domain-decomposed system");
    mg_connect(s, handles);
    SYS_CATCH(s)
    }

unsigned int b200_sys_get_n(b200_system* s) { return s->N; }

int b200_sys_set_tags(b200_system* s, const unsigned int* tags)
    {
    SYS_TRY(s)
    s->tag.upload(tags, s->N, s->stream);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    SYS_CATCH(s)
    }

int b200_sys_mg_stats(b200_system* s, unsigned long long* out8)
    {
    SYS_TRY(s)
    if (!s->mg) throw std::runtime_error("comm: not a domain-decomposed system");
    CUDA_OK(cudaMemcpyAsync(s->h_ms, s->d_ms.p, sizeof(MgState), cudaMemcpyDeviceToHost, s->stream));
    CUDA_OK(cudaStreamSynchronize(s->stream));
    out8[0] = s->N;
    out8[1] = s->Ncap;
    out8[2] = s->Gcap;
    out8[3] = s->h_ms->max_N;
    out8[4] = s->h_ms->max_send;
    out8[5] = s->h_ms->max_gsend;
    unsigned long long g = 0;
    for (int d = 0; d < MG_NDIR; d++) g += std::min(s->h_ms->gsend_cnt[d], s->mgc.seg_cap[d]);
    out8[6] = g;
    out8[7] = s->h_ms->epoch[0];
    SYS_CATCH(s)
    }

/* {wait ns, count} of the four barrier kinds (0 = MD step, 1..3 = rebuild phases), boundary chunks, chunks */
int b200_sys_mg_timing(b200_system* s, unsigned long long* out10)
    {
    SYS_TRY(s)
    if (!s->mg) throw std::runtime_error("comm: not a domain-decomposed system");
    CUDA_OK(cudaMemcpyAsync(s->h_ms, s->d_ms.p, sizeof(MgState), cudaMemcpyDeviceToHost, s->stream));
    CUDA_OK(cudaStreamSynchronize(s->stream));
    for (int k = 0; k < MG_NBARRIER; k++)
        {
        out10[2 * k] = s->h_ms->wait_ns[k];
        out10[2 * k + 1] = s->h_ms->wait_cnt[k];
        }
    out10[8] = s->h_ms->n_bchunks;
    out10[9] = s->mgl.nchunks;
    SYS_CATCH(s)
    }

/* ghost tags + positions as they currently sit in this rank's ghost region (validation) */
int b200_sys_mg_get_ghosts(b200_system* s, float* pos4, unsigned int* tag)
    {
    SYS_TRY(s)
    if (!s->mg) throw std::runtime_error("comm: not a domain-decomposed system");
    if (pos4)
        CUDA_OK(cudaMemcpyAsync(pos4, s->pos.p + s->Ncap, sizeof(float4) * s->Gcap, cudaMemcpyDeviceToHost, s->stream));
    if (tag)
        CUDA_OK(cudaMemcpyAsync(tag, s->tag.p + s->Ncap, sizeof(unsigned int) * s->Gcap, cudaMemcpyDeviceToHost, s->stream));
    CUDA_OK(cudaStreamSynchronize(s->stream));
    SYS_CATCH(s)
    }

void b200_sys_destroy(b200_system* s)
    {
    if (!s) return;
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    invalidate_graph(s);
    if (s->mg)
        {
        for (unsigned int q = 0; q < s->mgc.world; q++)
            if (q != s->mgc.rank && s->peer_base[q]) cudaIpcCloseMemHandle(s->peer_base[q]);
        if (s->slab) cudaFree(s->slab);
        if (s->h_ms) cudaFreeHost(s->h_ms);
        }
    if (s->h_flag) cudaFreeHost(s->h_flag);
    if (s->h_state) cudaFreeHost(s->h_state);
    cudaStream_t st = s->stream;
    delete s;
    if (st) cudaStreamDestroy(st);
    }

const char* b200_sys_last_error(b200_system* s) { return s->error.c_str(); }

int b200_sys_set_particles(b200_system* s, const float* pos4, const float* vel4, const int* image3)
    {
    SYS_TRY(s)
    if (pos4) s->pos.upload((const float4*)pos4, s->N, s->stream);
    if (vel4) s->vel.upload((const float4*)vel4, s->N, s->stream);
    if (image3) s->image.upload((const i3*)image3, s->N, s->stream);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    s->force_update = true;
    s->cl_force = true;
    SYS_CATCH(s)
    }

int b200_sys_set_accel(b200_system* s, const float* accel3)
    {
    SYS_TRY(s)
    s->accel.upload((const f3*)accel3, s->N, s->stream);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    s->accel_set = true;
    SYS_CATCH(s)
    }

int b200_sys_get_particles(b200_system* s, float* pos4, float* vel4, float* accel3, int* image3, unsigned int* tag,
                           unsigned int* rtag)
    {
    SYS_TRY(s)
    if (pos4) s->pos.download((float4*)pos4, s->N, s->stream);
    if (vel4) s->vel.download((float4*)vel4, s->N, s->stream);
    if (accel3) s->accel.download((f3*)accel3, s->N, s->stream);
    if (image3) s->image.download((i3*)image3, s->N, s->stream);
    if (tag) s->tag.download(tag, s->N, s->stream);
    if (rtag) s->rtag.download(rtag, s->N, s->stream);
    SYS_CATCH(s)
    }

int b200_sys_setup_pair(b200_system* s, int evaluator, float r_buff, unsigned int check_period, const float* rcut,
                        const float* params4, const float* ron, int shift_mode)
    {
    SYS_TRY(s)
    const unsigned int nt = s->ntypes;
    if (evaluator < B200_PAIR_LJ || evaluator > B200_PAIR_DPD) throw std::runtime_error("pair: unknown evaluator");
    if (evaluator != B200_PAIR_LJ && s->pair_mode == 1)
        throw std::runtime_error("pair: the cell-tile pair mode is instantiated for md.pair.lj only");
    s->pair_kind = evaluator;
    if (r_buff < 0.0f) throw std::runtime_error("nlist: Requested buffer radius is less than zero");
    if (shift_mode < 0 || shift_mode > 2) throw std::runtime_error("pair: invalid shift mode");
    s->r_buff = r_buff;
    s->every = check_period;
    s->shift_mode = shift_mode;
    s->h_r_cut.assign(rcut, rcut + nt * nt);
    s->d_r_cut.alloc(nt * nt); s->d_r_listsq.alloc(nt * nt); s->d_rcut_max.alloc(nt); s->d_maxsq.alloc(2 * (nt + 3));
    nlist_update_rlist(s);
    // pair tables: rcutsq = rcut*rcut, ronsq = ron*ron (PotentialPair::setRcut/setRon, hoomd/md/PotentialPair.h:237-269)
    std::vector<float2> prm(nt * nt);
    std::vector<float4> prm4(nt * nt);
    std::vector<float> rcsq(nt * nt), ronsq(nt * nt, 0.0f);
    for (unsigned int k = 0; k < nt * nt; k++)
        {
        prm[k] = make_float2(params4[4 * k], params4[4 * k + 1]);
        prm4[k] = make_float4(params4[4 * k], params4[4 * k + 1], params4[4 * k + 2], params4[4 * k + 3]);
        rcsq[k] = rcut[k] * rcut[k];
        if (ron) ronsq[k] = ron[k] * ron[k];
        }
    s->d_params.alloc(nt * nt); s->d_rcutsq.alloc(nt * nt); s->d_ronsq.alloc(nt * nt);
    s->d_params.upload(prm.data(), nt * nt, s->stream);
    s->d_params4.alloc(nt * nt);
    s->d_params4.upload(prm4.data(), nt * nt, s->stream);
    s->d_rcutsq.upload(rcsq.data(), nt * nt, s->stream);
    s->d_ronsq.upload(ronsq.data(), nt * nt, s->stream);
    // NeighborList ctor: Nmax = 8 per type, nlist 8*N (hoomd/md/NeighborList.cc:80-98)
    s->h_Nmax.assign(nt, 8u);
    s->d_Nmax.alloc(nt);
    s->d_Nmax.upload(s->h_Nmax.data(), nt, s->stream);
    s->nl_conditions.alloc(nt);
    s->head_list.alloc(s->Ncap);
    s->n_neigh.alloc(s->Ncap);
    s->nlist.alloc(8 * (size_t)s->Ncap);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    // CellList: nominal width = max r_cut + r_buff (NeighborListBinned.cc:56-76), radius 1, no ghosts (single domain)
    s->nominal_width = s->rcut_max_max + s->r_buff;
    if (s->mg && s->nominal_width > s->r_ghost * 1.000001f)
        throw std::runtime_error("comm: r_cut + r_buff exceeds the ghost layer width the domain was created with");
    b200_cell_list_dimensions(&s->box, s->nominal_width, 1, s->mg ? s->nominal_width : 0.0f, &s->dim, s->ghost_width);
    const size_t ncells = (size_t)s->dim.x * s->dim.y * s->dim.z;
    s->cell_count.alloc(ncells); s->cell_start.alloc(ncells + 1);
    const size_t ntot = (size_t)s->Ncap + s->Gcap;
    s->cell_xyzf.alloc(ntot); s->cell_type.alloc(ntot);
    s->particle_cell.alloc(ntot); s->particle_slot.alloc(ntot); s->scratch_idx.alloc(ntot);
    s->have_nlist = true;
    s->force_update = true;
    s->cl_force = true;
    s->nl_first_compute = true;
    for (int d = 0; d < 3; d++) s->last_L[d] = s->box.L[d]; // NeighborList ctor (NeighborList.cc:58)
    invalidate_graph(s);
    tile_configure(s);
    SYS_CATCH(s)
    }

int b200_sys_setup_lj(b200_system* s, float r_buff, unsigned int check_period, const float* rcut, const float* lj1,
                      const float* lj2, const float* ron, int shift_mode)
    {
    const unsigned int nt = s->ntypes;
    std::vector<float> p4(4 * (size_t)nt * nt, 0.0f);
    for (unsigned int k = 0; k < nt * nt; k++)
        {
        p4[4 * k] = lj1[k];
        p4[4 * k + 1] = lj2[k];
        }
    return b200_sys_setup_pair(s, B200_PAIR_LJ, r_buff, check_period, rcut, p4.data(), ron, shift_mode);
    }

int b200_sys_set_bonds(b200_system* s, unsigned int nbonds, const unsigned int* tags2, const unsigned int* types,
                       unsigned int n_bond_types)
    {
    SYS_TRY(s)
    // domain decomposition: the table covers the GLOBAL tags on every rank and stays keyed by tag (a partner on another rank
    // is reached through its ghost copy, comm.cu:k_mg_rtag_fill)
    const unsigned int N = s->mg ? s->mg_n_global : s->N;
    if (s->mg)
        {
        s->idx_tables = false;
        if (s->rtag.n < N) s->rtag.alloc(N);
        }
    if (nbonds && n_bond_types == 0) throw std::runtime_error("bond: no bond types defined");
    if (n_bond_types > 256) throw std::runtime_error("bond: more than 256 bond types");
    std::vector<unsigned int> cnt(N, 0u);
    for (unsigned int b = 0; b < nbonds; b++)
        {
        const unsigned int ta = tags2[2 * b], tb = tags2[2 * b + 1];
        if (ta >= N || tb >= N) throw std::runtime_error("bond: particle tag out of bounds when attempting to add a bond");
        if (ta == tb) throw std::runtime_error("bond: particle cannot be bonded to itself");
        if (types[b] >= n_bond_types) throw std::runtime_error("bond: invalid bond type");
        cnt[ta]++; cnt[tb]++;
        }
    unsigned int mx = 0;
    for (unsigned int i = 0; i < N; i++) mx = std::max(mx, cnt[i]);
    s->h_bond_tags.assign(tags2, tags2 + 2 * (size_t)nbonds);
    s->h_bond_types.assign(types, types + nbonds);
    s->n_bond_types = n_bond_types;
    s->bond_max = mx;
    s->bond_pitch = N;
    // per-particle rows in bond order, like BondedGroupData::rebuildGPUTable fills its table (hoomd/BondedGroupData.cc)
    std::vector<uint2> table((size_t)std::max(mx, 1u) * N, make_uint2(0u, 0u));
    std::fill(cnt.begin(), cnt.end(), 0u);
    for (unsigned int b = 0; b < nbonds; b++)
        {
        const unsigned int ta = tags2[2 * b], tb = tags2[2 * b + 1];
        table[(size_t)cnt[ta]++ * N + ta] = make_uint2(tb, types[b]);
        table[(size_t)cnt[tb]++ * N + tb] = make_uint2(ta, types[b]);
        }
    s->bond_n.alloc(N);
    s->bond_list.alloc(table.size());
    s->bond_n.upload(cnt.data(), N, s->stream);
    s->bond_list.upload(table.data(), table.size(), s->stream);
    s->bond_n_idx.alloc(N);
    s->bond_list_idx.alloc(table.size());
    s->idx_tables_dirty = true;
    CUDA_OK(cudaStreamSynchronize(s->stream));
    invalidate_graph(s);
    SYS_CATCH(s)
    }

int b200_sys_setup_bond_force(b200_system* s, int kind, const float* K, const float* r0, const float* lj1,
                              const float* lj2)
    {
    SYS_TRY(s)
    if (kind != 0 && kind != 1) throw std::runtime_error("bond: unknown bond potential");
    if (s->n_bond_types == 0 || s->h_bond_types.empty()) throw std::runtime_error("bond: No bonds are defined.");
    if (s->tile_on) throw std::runtime_error("bond: the cell-tile pair path does not carry bonded forces (use pair mode 0)");
    const unsigned int nb = s->n_bond_types;
    if (kind == 0)
        {
        std::vector<float4> p(nb);
        for (unsigned int t = 0; t < nb; t++) p[t] = make_float4(K[t], r0[t], lj1[t], lj2[t]);
        s->bond_params4.alloc(nb);
        s->bond_params4.upload(p.data(), nb, s->stream);
        }
    else
        {
        std::vector<float2> p(nb);
        for (unsigned int t = 0; t < nb; t++) p[t] = make_float2(K[t], r0[t]);
        s->bond_params2.alloc(nb);
        s->bond_params2.upload(p.data(), nb, s->stream);
        }
    if (s->mg && kind == 0)
        for (unsigned int t = 0; t < nb; t++)
            if (r0[t] > s->r_ghost)
                throw std::runtime_error("bond.fene: r0 exceeds the ghost layer width (r_cut + r_buff) of the decomposition");
    s->bond_force.alloc(s->Ncap);
    s->pair_force_copy.alloc(s->Ncap);
    s->bond_virial.alloc(6 * (size_t)s->virial_pitch);
    s->pair_virial_copy.alloc(6 * (size_t)s->virial_pitch);
    s->bond_flags.alloc(1);
    CUDA_OK(cudaMemsetAsync(s->bond_flags.p, 0, sizeof(unsigned int), s->stream));
    CUDA_OK(cudaStreamSynchronize(s->stream));
    s->bond_kind = kind;
    invalidate_graph(s);
    SYS_CATCH(s)
    }

int b200_sys_nlist_exclusions_from_bonds(b200_system* s, int on)
    {
    SYS_TRY(s)
    if (s->tile_on) throw std::runtime_error("nlist: the cell-tile pair path does not filter exclusions (use pair mode 0)");
    invalidate_graph(s);
    s->force_update = true;
    if (!on)
        {
        s->exclusions_set = false; // NeighborList::clearExclusions (hoomd/md/NeighborList.cc:505-520)
        s->ex_max = 0;
        }
    else
        {
        const unsigned int N = s->mg ? s->mg_n_global : s->N;
        if (s->mg)
            {
            s->idx_tables = false;
            if (s->rtag.n < N) s->rtag.alloc(N);
            }
        const size_t nbonds = s->h_bond_types.size();
        // NeighborList::addExclusion per bond, duplicates skipped (hoomd/md/NeighborList.cc:431-470)
        std::vector<std::vector<unsigned int>> ex(N);
        for (size_t b = 0; b < nbonds; b++)
            {
            const unsigned int ta = s->h_bond_tags[2 * b], tb = s->h_bond_tags[2 * b + 1];
            if (std::find(ex[ta].begin(), ex[ta].end(), tb) != ex[ta].end()) continue;
            ex[ta].push_back(tb);
            ex[tb].push_back(ta);
            }
        unsigned int mx = 0;
        std::vector<unsigned int> cnt(N);
        for (unsigned int i = 0; i < N; i++)
            {
            cnt[i] = (unsigned int)ex[i].size();
            mx = std::max(mx, cnt[i]);
            }
        std::vector<unsigned int> table((size_t)std::max(mx, 1u) * N, 0u);
        for (unsigned int i = 0; i < N; i++)
            for (unsigned int k = 0; k < cnt[i]; k++) table[(size_t)k * N + i] = ex[i][k];
        s->ex_n.alloc(N);
        s->ex_list.alloc(table.size());
        s->ex_n.upload(cnt.data(), N, s->stream);
        s->ex_list.upload(table.data(), table.size(), s->stream);
        s->ex_n_idx.alloc(N);
        s->ex_list_idx.alloc(table.size());
        s->idx_tables_dirty = true;
        CUDA_OK(cudaStreamSynchronize(s->stream));
        s->ex_max = mx;
        s->ex_pitch = N;
        s->exclusions_set = true; // addExclusion sets it even when the bond list is empty
        }
    SYS_CATCH(s)
    }

int b200_sys_get_force_component(b200_system* s, int which, float* force4, float* virial6)
    {
    SYS_TRY(s)
    const unsigned int N = s->N;
    const float4* f = nullptr;
    const float* v = nullptr;
    if (s->bond_kind < 0)
        {
        if (which != 0) throw std::runtime_error("force: no bond force set up");
        f = s->net_force.p; v = s->net_virial.p;
        }
    else if (which == 0) { f = s->pair_force_copy.p; v = s->pair_virial_copy.p; }
    else if (which == 1) { f = s->bond_force.p; v = s->bond_virial.p; }
    else throw std::runtime_error("force: unknown component");
    if (force4) CUDA_OK(cudaMemcpyAsync(force4, f, sizeof(float4) * N, cudaMemcpyDeviceToHost, s->stream));
    if (virial6)
        for (int c = 0; c < 6; c++)
            CUDA_OK(cudaMemcpyAsync(virial6 + (size_t)c * N, v + (size_t)c * s->virial_pitch, sizeof(float) * N,
                                    cudaMemcpyDeviceToHost, s->stream));
    CUDA_OK(cudaStreamSynchronize(s->stream));
    SYS_CATCH(s)
    }

int b200_sys_set_virial(b200_system* s, int on)
    {
    s->compute_virial = on != 0;
    invalidate_graph(s);
    return 0;
    }

int b200_sys_setup_integrator(b200_system* s, float dt)
    {
    s->dt = dt;
    s->have_integrator = true;
    invalidate_graph(s);
    return 0;
    }

int b200_sys_add_nve(b200_system* s, int limit, float limit_val, int zero_force)
    {
    s->method = 1;
    s->limit = limit;
    s->limit_val = limit_val;
    s->zero_force = zero_force;
    invalidate_graph(s);
    return 0;
    }

int b200_sys_add_langevin(b200_system* s, float kT, unsigned int seed, const float* gamma, int noiseless_t)
    {
    SYS_TRY(s)
    s->method = 2;
    s->limit = 0;
    s->zero_force = 0;
    s->kT = kT;
    s->seed_hashed = b200_langevin_hash_seed(seed);
    s->noiseless_t = noiseless_t;
    std::vector<float> g(s->ntypes, 1.0f);
    if (gamma) g.assign(gamma, gamma + s->ntypes);
    s->d_gamma.alloc(s->ntypes);
    s->d_gamma.upload(g.data(), s->ntypes, s->stream);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    invalidate_graph(s);
    SYS_CATCH(s)
    }

// md.integrate.langevin(..., dscale=lambda, tally=...) (hoomd/md/integrate.py:844-940; TwoStepLangevinBase use_lambda / lambda,
// TwoStepLangevin::setTally).  Call after b200_sys_add_langevin.
int b200_sys_set_langevin_options(b200_system* s, int use_lambda, float lambda, int tally)
    {
    SYS_TRY(s)
    if (s->method != 2) throw std::runtime_error("integrate.langevin: no Langevin method set");
    if (s->mg && (use_lambda || tally)) throw std::runtime_error("integrate.langevin: dscale / tally are not available under domain decomposition yet");
    if (s->tile_on && (use_lambda || tally)) throw std::runtime_error("integrate.langevin: dscale / tally are not available on the cell-tile path");
    s->use_lambda = use_lambda != 0;
    s->lambda = lambda;
    s->tally = tally != 0;
    if (s->use_lambda && !s->diameter.n)
        {
        std::vector<float> one(s->N, 1.0f); // ParticleData default diameter
        s->diameter.alloc(s->N);
        s->diameter.upload(one.data(), s->N, s->stream);
        }
    if (!s->tally_acc.n)
        {
        s->tally_acc.alloc(3);
        s->ck_tally.alloc(3);
        s->tally_tmp.alloc(1);
        }
    CUDA_OK(cudaMemsetAsync(s->tally_acc.p, 0, sizeof(double) * 3, s->stream)); // m_reservoir_energy(0), m_extra_energy_overdeltaT(0)
    CUDA_OK(cudaStreamSynchronize(s->stream));
    invalidate_graph(s);
    SYS_CATCH(s)
    }

// per-particle diameters by TAG (ParticleData::setDiameter); only md.integrate.langevin(dscale=...) reads them on this path
int b200_sys_set_diameters(b200_system* s, const float* diameter_by_tag)
    {
    SYS_TRY(s)
    if (s->mg) throw std::runtime_error("particle diameters are not carried under domain decomposition yet");
    if (!s->diameter.n) s->diameter.alloc(s->N);
    s->diameter.upload(diameter_by_tag, s->N, s->stream);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    invalidate_graph(s);
    SYS_CATCH(s)
    }

// the "langevin_reservoir_energy" log quantity (TwoStepLangevin::getLogValue, hoomd/md/TwoStepLangevin.cc:71-80):
// m_reservoir_energy + m_extra_energy_overdeltaT * dt with m_reservoir_energy = -dt * sum over steps of E_step and
// m_extra_energy_overdeltaT = E_last / 2 (:416-417)
int b200_sys_get_langevin_reservoir_energy(b200_system* s, double* out)
    {
    SYS_TRY(s)
    if (s->method != 2 || !s->tally) throw std::runtime_error("integrate.langevin: tally is not enabled");
    double h[3];
    CUDA_OK(cudaStreamSynchronize(s->stream));
    CUDA_OK(cudaMemcpy(h, s->tally_acc.p, sizeof(h), cudaMemcpyDeviceToHost));
    const double e_last = s->timestep ? h[1u + ((s->timestep - 1u) & 1u)] : 0.0;
    *out = (double)(float)(-(double)s->dt * h[0] + 0.5 * e_last * (double)s->dt);
    SYS_CATCH(s)
    }

// hoomd.group.*: the particles the integration method acts on, as member TAGS (ParticleGroup, hoomd/ParticleGroup.h); n = 0
// restores group.all().  Non-members keep their position and velocity; forces on them are still computed and reported.
int b200_sys_set_group(b200_system* s, const unsigned int* member_tags, unsigned int n)
    {
    SYS_TRY(s)
    if (s->mg) throw std::runtime_error("group: integration groups other than all are not available under domain decomposition yet");
    if (s->tile_on) throw std::runtime_error("group: integration groups other than all are not available on the cell-tile path");
    if (n == 0 || !member_tags)
        {
        s->group_flag.release();
        s->group_size = 0;
        }
    else
        {
        std::vector<unsigned char> f(s->N, 0);
        unsigned int cnt = 0;
        for (unsigned int k = 0; k < n; k++)
            {
            if (member_tags[k] >= s->N) throw std::runtime_error("group: particle tag out of bounds");
            if (!f[member_tags[k]]) cnt++;
            f[member_tags[k]] = 1;
            }
        s->group_flag.alloc(s->N);
        s->group_flag.upload(f.data(), s->N, s->stream);
        s->group_size = cnt;
        CUDA_OK(cudaStreamSynchronize(s->stream));
        }
    s->group_idx_dirty = true;
    if (s->method == 3) // NDOF of the thermostat follows the group (IntegratorTwoStep::getNDOF)
        s->nvt_scale = std::exp2(std::floor(42.0 - std::log2((double)nvt_ndof(s) * (double)s->kT)));
    invalidate_graph(s);
    SYS_CATCH(s)
    }

// hoomd.update.box_resize: BoxResizeUpdater::update (hoomd/BoxResizeUpdater.cc:60-136) for one target box.  The
// reference's `no_change` test divides the (zero) tilt difference by the zero tilt: NaN < 1e-7 is false, so an orthorhombic
// box is ALWAYS re-set, even to the same lengths -- reproduced here (the cell list is forced, positions are re-scaled).
int b200_sys_set_box(b200_system* s, float Lx, float Ly, float Lz, int scale_particles)
    {
    SYS_TRY(s)
    if (s->mg) throw std::runtime_error("update.box_resize: not available under domain decomposition yet");
    if (s->tile_on) throw std::runtime_error("update.box_resize: not available on the cell-tile path");
    if (!(Lx > 0.0f && Ly > 0.0f && Lz > 0.0f)) throw std::runtime_error("update.box_resize: box lengths must be positive");
    // updaters run in the order they were added: the sorter (created with the system) before box_resize
    if (s->sorter_period && s->timestep == s->sorter_next)
        {
        sorter_update(s);
        s->sorter_next += s->sorter_period;
        }
    b200_box nb;
    b200_box_init(&nb, Lx, Ly, Lz);
    for (int d = 0; d < 3; d++) nb.periodic[d] = s->box.periodic[d];
    CUDA_OK(b200_box_resize_particles((float*)s->pos.p, (int*)s->image.p, s->N, &s->box, &nb, scale_particles, s->stream));
    s->kernels_last_run += 1;
    apply_new_box(s, nb);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    SYS_CATCH(s)
    }

void b200_sys_get_box(b200_system* s, float* L3)
    {
    for (int d = 0; d < 3; d++) L3[d] = s->box.L[d];
    }

// method.set_params(kT=...) / a temperature variant evaluated by the caller once per step: TwoStepLangevinBase::setT,
// TwoStepNVTMTK::setT.  The set point is a by-value argument of the captured kernels, so the step graph is invalidated; chunks
// of <= 16 steps behind an invalid graph launch their kernels directly (run_chunk_device).
int b200_sys_set_temperature(b200_system* s, float kT)
    {
    SYS_TRY(s)
    if (s->method != 2 && s->method != 3 && s->method != 4)
        throw std::runtime_error("integrate: the active method has no temperature");
    if (s->method >= 3 && !(kT > 0.0f)) throw std::runtime_error("integrate.nvt / npt: kT must be positive");
    if (kT != s->kT)
        {
        s->kT = kT;
        invalidate_graph(s);
        }
    SYS_CATCH(s)
    }

// md.nlist.cell(dist_check=...) / nlist.set_params(dist_check=...): NeighborList::setEvery(every, dist_check)
int b200_sys_set_dist_check(b200_system* s, int on)
    {
    SYS_TRY(s)
    if (s->mg && !on) throw std::runtime_error("nlist: dist_check=False is not available under domain decomposition yet");
    s->dist_check = on != 0;
    s->force_update = true; // NeighborList::setEvery ends in forceUpdate() (hoomd/md/NeighborList.h:148-153)
    SYS_CATCH(s)
    }

int b200_sys_add_nvt(b200_system* s, float kT, float tau)
    {
    SYS_TRY(s)
    if (s->mg && s->group_flag.n) throw std::runtime_error("integrate.nvt: groups are not available under domain decomposition");
    if (s->tile_on) throw std::runtime_error("integrate.nvt: not available on the cell-tile path");
    if (!(kT > 0.0f)) throw std::runtime_error("integrate.nvt: kT must be positive");
    // under decomposition the kinetic energy is summed over the ranks by the stand-alone step barrier (comm.cu)
    if (s->mg) s->mg_mode &= ~4;
    if (tau <= 0.0f) fprintf(stderr, "*Warning*: integrate.nvt: tau set less than 0.0\n"); // TwoStepNVTMTK.cc:46-47
    s->method = 3;
    s->limit = 0;
    s->zero_force = 0;
    s->kT = kT;
    s->nvt_tau = tau;
    // initializeIntegratorVariables: xi = eta = 0, m_exp_thermo_fac = 1 (TwoStepNVTMTK.cc:39,50-54); gen = "no advance yet"
    float init[4] = {0.0f, 0.0f, 1.0f, 0.0f};
    const unsigned int no_gen = 0xfffffffeu;
    memcpy(&init[3], &no_gen, 4);
    s->d_nvt.alloc(4);
    s->ck_nvt.alloc(4);
    s->d_nvt.upload(init, 4, s->stream);
    s->nvt_partial.alloc(b200_system::NVT_BLOCKS);
    // fixed-point accumulators of sum(m v^2): 2^62 units span 2^20 x the set point's ndof kT (a power-of-two scale, so the
    // conversion only rounds to an integer: 2^-42 of the expected sum per contribution)
    s->nvt_acc.alloc(2 * NVT_SLOTS * NVT_SLOT_STRIDE);
    CUDA_OK(cudaMemsetAsync(s->nvt_acc.p, 0, sizeof(unsigned long long) * s->nvt_acc.n, s->stream));
    s->nvt_scale = std::exp2(std::floor(42.0 - std::log2((double)nvt_ndof(s) * (double)kT)));
    CUDA_OK(cudaStreamSynchronize(s->stream));
    invalidate_graph(s);
    SYS_CATCH(s)
    }

// Re-configuration within one job: the IntegratorVariables of the engine this one replaces.  m_exp_thermo_fac is a pure
// function of xi (TwoStepNVTMTK.cc:402), so a continuing run keeps the factor the last advanceThermostat left behind.
int b200_sys_set_nvt_state(b200_system* s, const float* xi_eta)
    {
    SYS_TRY(s)
    if (s->method != 3) throw std::runtime_error("integrate.nvt: no NVT method set");
    const float arg = -0.5f * xi_eta[0] * s->dt;
    float h[4] = {xi_eta[0], xi_eta[1], (float)std::exp((double)arg), 0.0f};
    const unsigned int no_gen = 0xfffffffeu;
    memcpy(&h[3], &no_gen, 4);
    s->d_nvt.upload(h, 4, s->stream);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    SYS_CATCH(s)
    }

int b200_sys_get_nvt_state(b200_system* s, float* xi_eta)
    {
    SYS_TRY(s)
    if (s->method != 3) throw std::runtime_error("integrate.nvt: no NVT method set");
    float h[4];
    CUDA_OK(cudaStreamSynchronize(s->stream));
    CUDA_OK(cudaMemcpy(h, s->d_nvt.p, sizeof(h), cudaMemcpyDeviceToHost));
    xi_eta[0] = h[0];
    xi_eta[1] = h[1];
    SYS_CATCH(s)
    }

int b200_sys_set_dpd_thermostat(b200_system* s, float kT, unsigned int seed)
    {
    SYS_TRY(s)
    if (kT > 0.0f)
        {
        if (!s->have_nlist || s->pair_kind != B200_PAIR_DPD)
            throw std::runtime_error("pair.dpd: set up the DPD pair force first (b200_sys_setup_pair with B200_PAIR_DPD)");
        if (s->mg) throw std::runtime_error("pair.dpd: not available under domain decomposition yet");
        if (s->tile_on) throw std::runtime_error("pair.dpd: not available on the cell-tile path");
        }
    s->dpd_thermo = kT > 0.0f;
    s->dpd_kT = kT;
    s->dpd_seed = b200_langevin_hash_seed(seed); // PotentialPairDPDThermo::setSeed hashes the user seed the same way
    invalidate_graph(s);
    SYS_CATCH(s)
    }

// md.integrate.npt(group, kT, tau, P, tauP, couple, x, y, z, nph, rescale_all, gamma) (hoomd/md/integrate.py:406-506)
int b200_sys_add_npt(b200_system* s, float kT, float tau, float P, float tauP, int couple, unsigned int flags, int nph,
                     int rescale_all, float gamma)
    {
    SYS_TRY(s)
    if (s->mg) throw std::runtime_error("integrate.npt: not available under domain decomposition yet");
    if (s->tile_on) throw std::runtime_error("integrate.npt: not available on the cell-tile path");
    if (flags & ~7u) throw std::runtime_error("integrate.npt: tilt degrees of freedom (xy, xz, yz) need triclinic boxes");
    if (couple < 0 || couple > 4) throw std::runtime_error("integrate.npt: Invalid coupling mode.");
    if (!nph && !(kT > 0.0f)) throw std::runtime_error("integrate.npt: kT must be positive");
    if (tau <= 0.0f) fprintf(stderr, "*Warning*: integrate.npt: tau set less than 0.0\n");   // TwoStepNPTMTK.cc:67-70
    if (tauP <= 0.0f) fprintf(stderr, "*Warning*: integrate.npt: tauP set less than 0.0\n");
    if (flags == 0) fprintf(stderr, "*Warning*: integrate.npt: No barostat couplings specified.\n");
    s->method = 4;
    s->limit = 0;
    s->zero_force = 0;
    s->kT = kT;
    s->npt = b200_system::Npt();
    s->npt.tau = tau;
    s->npt.tauP = tauP;
    s->npt.P = P;
    s->npt.couple = couple;
    s->npt.flags = flags;
    s->npt.nph = nph != 0;
    s->npt.rescale_all = rescale_all != 0;
    s->npt.gamma = gamma;
    // the method requests the pressure tensor on every step (TwoStepNPTMTK.h:152-163)
    s->compute_virial = 1;
    invalidate_graph(s);
    SYS_CATCH(s)
    }

/* out12: the method's IntegratorVariables {eta, xi, nuxx, nuxy, nuxz, nuyy, nuyz, nuzz, xi_rot, eta_rot} and its log
 * quantities {npt_thermostat_energy, npt_barostat_energy} (TwoStepNPTMTK.cc:593-636) */
int b200_sys_get_npt_state(b200_system* s, float* out12)
    {
    SYS_TRY(s)
    if (s->method != 4) throw std::runtime_error("integrate.npt: no NPT method set");
    const b200_system::Npt& n = s->npt;
    const float v[10] = {n.eta, n.xi, n.nuxx, 0.0f, 0.0f, n.nuyy, 0.0f, n.nuzz, 0.0f, 0.0f};
    for (int k = 0; k < 10; k++) out12[k] = v[k];
    const unsigned int ndof = nvt_ndof(s), d = 3;
    const double T = (double)s->kT;
    out12[10] = (float)((double)ndof * T * (double)(n.eta + n.tau * n.tau * n.xi * n.xi / 2.0f));
    const float W = (float)((double)((float)(n.ndof + d) / (float)d) * T * n.tauP * n.tauP);
    out12[11] = W * (n.nuxx * n.nuxx + n.nuyy * n.nuyy + n.nuzz * n.nuzz + 0.0f + 0.0f + 0.0f) / 2.0f;
    SYS_CATCH(s)
    }

int b200_sys_set_npt_state(b200_system* s, const float* vars10)
    {
    SYS_TRY(s)
    if (s->method != 4) throw std::runtime_error("integrate.npt: no NPT method set");
    s->npt.eta = vars10[0];
    s->npt.xi = vars10[1];
    s->npt.nuxx = vars10[2];
    s->npt.nuyy = vars10[5];
    s->npt.nuzz = vars10[7];
    SYS_CATCH(s)
    }

/* ComputeThermo::getPressureTensor of the integration group at the current step: {xx, xy, xz, yy, yz, zz, P, KE} */
int b200_sys_get_pressure_tensor(b200_system* s, float* out8)
    {
    SYS_TRY(s)
    if (!s->compute_virial) throw std::runtime_error("compute.thermo: the pressure tensor needs the virial (b200_sys_set_virial)");
    float P6[6], ke;
    npt_thermo(s, true, P6, &ke);
    for (int k = 0; k < 6; k++) out8[k] = P6[k];
    // ComputeThermo.cc:352: isotropic pressure from the same sums (W = trace of the virial / 3)
    double h[13];
    s->npt_out.download(h, 13, s->stream);
    const float volume = s->box.L[0] * s->box.L[1] * s->box.L[2];
    const double ke_d = 0.5f * (h[0] + h[3] + h[5]);
    const double W = (float)(1. / 3.) * (h[7] + h[10] + h[12]);
    out8[6] = (float)((2.0 * ke_d / 3.0f + W) / volume);
    out8[7] = ke;
    SYS_CATCH(s)
    }

int b200_sys_set_sorter(b200_system* s, unsigned int period)
    {
    s->sorter_period = period;
    s->sorter_next = s->timestep;
    return 0;
    }

int b200_sys_sort_now(b200_system* s)
    {
    SYS_TRY(s)
    sorter_update(s, false);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    SYS_CATCH(s)
    }

void b200_sys_set_timestep(b200_system* s, unsigned int ts) { s->timestep = ts; }
unsigned int b200_sys_get_timestep(b200_system* s) { return s->timestep; }

int b200_sys_set_tuning(b200_system* s, unsigned int pair_tpp, int use_graph)
    {
    s->pair_tpp = pair_tpp;
    s->use_graph = use_graph;
    invalidate_graph(s);
    return 0;
    }

int b200_sys_set_pair_mode(b200_system* s, int mode, unsigned int bx, unsigned int by, unsigned int bz, unsigned int tpp)
    {
    SYS_TRY(s)
    if (mode < -1 || mode > 1) throw std::runtime_error("pair_mode must be -1 (auto), 0 (CSR walk) or 1 (cell tiles)");
    if (tpp != 0 && tpp != 1 && tpp != 2 && tpp != 4) throw std::runtime_error("tile: tpp must be 0 (auto), 1, 2 or 4");
    s->pair_mode = mode;
    s->tile_req[0] = bx; s->tile_req[1] = by; s->tile_req[2] = bz;
    if (bx == 0 || by == 0 || bz == 0) s->tile_req[0] = s->tile_req[1] = s->tile_req[2] = 0;
    s->tile_tpp = tpp;
    const bool was = s->tile_on;
    tile_configure(s);
    if (was != s->tile_on || s->tile_on)
        {
        s->force_update = true;
        s->cl_force = true;
        }
    SYS_CATCH(s)
    }

int b200_sys_get_pair_mode(b200_system* s, unsigned int* out8)
    {
    out8[0] = s->tile_on ? 1u : 0u;
    out8[1] = s->tile_geom.b.x; out8[2] = s->tile_geom.b.y; out8[3] = s->tile_geom.b.z;
    out8[4] = s->tile_geom.cap; out8[5] = s->tile_tpp; out8[6] = s->tile_Nmax;
    out8[7] = s->tile_geom.tdim.x * s->tile_geom.tdim.y * s->tile_geom.tdim.z;
    return 0;
    }

/* out4: {stepping mode (b200_sys_set_tuning's use_graph), MD steps per graph unit (0 = data dependent: multi-step kernel),
 * 1 when the multi-step kernel k_md_steps drives the graph, graph units launched by the last run} */
int b200_sys_get_stepping(b200_system* s, unsigned int* out4)
    {
    out4[0] = (unsigned int)s->use_graph;
    out4[1] = (s->graph_valid && s->graph_persist) ? 0u : s->graph_steps;
    out4[2] = (s->graph_valid && s->graph_persist) ? 1u : 0u;
    out4[3] = (unsigned int)std::min<unsigned long long>(s->persist_units_last_run, 0xffffffffull);
    return 0;
    }

/* multi-step launches (k_md_steps): -1 automatic (default), 0 off.  Takes effect at the next graph build. */
int b200_sys_set_multistep(b200_system* s, int mode)
    {
    s->persist_mode = mode ? -1 : 0;
    invalidate_graph(s);
    return 0;
    }

int b200_sys_compute_nlist(b200_system* s, unsigned int ts)
    {
    SYS_TRY(s)
    nlist_compute(s, ts);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    SYS_CATCH(s)
    }

int b200_sys_compute_forces(b200_system* s, unsigned int ts)
    {
    SYS_TRY(s)
    if (s->mg)
        mg_prep_run(s, ts);
    else
        compute_net_force(s, ts);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    SYS_CATCH(s)
    }

int b200_sys_run(b200_system* s, unsigned int nsteps, double* elapsed_s, double* gpu_ms)
    {
    SYS_TRY(s)
    s->kernels_last_run = 0;
    if (s->mg)
        {
        if (!s->have_integrator || !s->method) throw std::runtime_error("integrate: no integration method set");
        if (s->use_graph < 2) throw std::runtime_error("comm: domain decomposition needs device-driven stepping");
        mg_prep_run(s, s->timestep);
        }
    else
        prep_run(s, s->timestep);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    EventPair ev;
    const cudaEvent_t e0 = ev.e0, e1 = ev.e1;
    auto t0 = std::chrono::steady_clock::now();
    CUDA_OK(cudaEventRecord(e0, s->stream));
    unsigned int remaining = nsteps;
    const unsigned int chunk_max = 500;
    while (remaining)
        {
        if (s->sorter_period && s->timestep == s->sorter_next)
            {
            sorter_update(s);
            s->sorter_next += s->sorter_period;
            }
        unsigned int chunk = std::min(remaining, chunk_max);
        if (s->sorter_period && s->sorter_next > s->timestep) chunk = std::min(chunk, s->sorter_next - s->timestep);
        if (s->mg)
            mg_run_chunk(s, chunk);
        else if (s->use_graph && s->method != 4 && !s->dpd_thermo)
            run_chunk_device(s, chunk);
        else
            for (unsigned int k = 0; k < chunk; k++) host_step(s);
        remaining -= chunk;
        }
    CUDA_OK(cudaEventRecord(e1, s->stream));
    CUDA_OK(cudaStreamSynchronize(s->stream));
    auto t1 = std::chrono::steady_clock::now();
    float ms = 0.0f;
    CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
    if (elapsed_s) *elapsed_s = std::chrono::duration<double>(t1 - t0).count();
    if (gpu_ms) *gpu_ms = ms;
    SYS_CATCH(s)
    }

int b200_sys_step_one(b200_system* s, unsigned int ts)
    {
    SYS_TRY(s)
    if (!s->method) throw std::runtime_error("integrate: no integration method set");
    integrate_step_one(s, ts);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    SYS_CATCH(s)
    }

int b200_sys_step_two(b200_system* s, unsigned int ts, const float* net_force4)
    {
    SYS_TRY(s)
    if (!s->method) throw std::runtime_error("integrate: no integration method set");
    if (net_force4) s->net_force.upload((const float4*)net_force4, s->N, s->stream);
    integrate_step_two(s, ts);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    SYS_CATCH(s)
    }

int b200_sys_get_forces(b200_system* s, float* force4, float* virial6)
    {
    SYS_TRY(s)
    if (force4) s->net_force.download((float4*)force4, s->N, s->stream);
    if (virial6)
        {
        for (unsigned int k = 0; k < 6; k++)
            CUDA_OK(cudaMemcpyAsync(virial6 + (size_t)k * s->N, s->net_virial.p + (size_t)k * s->virial_pitch,
                                    sizeof(float) * s->N, cudaMemcpyDeviceToHost, s->stream));
        CUDA_OK(cudaStreamSynchronize(s->stream));
        }
    SYS_CATCH(s)
    }

unsigned int b200_sys_get_nlist_size(b200_system* s)
    {
    return s->tile_on ? s->N * s->tile_Nmax : (unsigned int)s->nlist.n;
    }

int b200_sys_get_nlist(b200_system* s, unsigned int* n_neigh, unsigned int* head, unsigned int* nlist,
                       unsigned int nlist_len)
    {
    SYS_TRY(s)
    if (s->tile_on)
        {
        // tile slots -> particle indices (current storage order), rows of tile_Nmax entries
        DevBuf<unsigned int> on, oh, ol;
        on.alloc(s->N); oh.alloc(s->N); ol.alloc((size_t)s->N * s->tile_Nmax);
        TileBuildArgs ba;
        tile_build_args(s, ba, nullptr, s->pos.p);
        CUDA_OK(launch_tile_to_csr(ba, on.p, oh.p, ol.p, s->stream));
        if (n_neigh) on.download(n_neigh, s->N, s->stream);
        if (head) oh.download(head, s->N, s->stream);
        if (nlist) ol.download(nlist, std::min<size_t>(nlist_len, ol.n), s->stream);
        CUDA_OK(cudaStreamSynchronize(s->stream));
        return 0;
        }
    if (n_neigh) s->n_neigh.download(n_neigh, s->N, s->stream);
    if (head) s->head_list.download(head, s->N, s->stream);
    if (nlist) s->nlist.download(nlist, std::min<size_t>(nlist_len, s->nlist.n), s->stream);
    SYS_CATCH(s)
    }

int b200_sys_get_stats(b200_system* s, unsigned long long* out8)
    {
    out8[0] = s->num_updates;
    out8[1] = s->num_forced;
    out8[2] = s->num_overflows;
    out8[3] = s->kernels_last_run;
    out8[4] = s->dim.x;
    out8[5] = s->dim.y;
    out8[6] = s->dim.z;
    out8[7] = s->tile_on ? s->tile_Nmax : (s->h_Nmax.empty() ? 0 : s->h_Nmax[0]);
    return 0;
    }

int b200_sys_time_kernel(b200_system* s, int which, unsigned int iters, double* avg_ms)
    {
    SYS_TRY(s)
    if (!s->have_nlist) throw std::runtime_error("nlist: not set up");
    if (iters == 0) iters = 1;
    ensure_idx_tables(s);
    EventPair ev;
    const cudaEvent_t e0 = ev.e0, e1 = ev.e1;
    b200_pair_args pa;
    pair_args(s, pa);
    b200_cell_list_args ca;
    cell_list_args(s, ca);
    b200_nlist_args na;
    nlist_args(s, na);
    if (which == 2 || which == 5)
        {
        // the single-kernel MD step itself (5: `iters` steps inside ONE launch of the multi-step kernel k_md_steps): `iters` back-to-back launches (alternating ping-pong phase, no rebuild in
        // between, never the chunk's last step) on a scratch copy of the integrator state, which is restored afterwards
        if (!s->method) throw std::runtime_error("integrate: no integration method set");
        const unsigned int N = s->N;
        if (s->ck_pos.n < N)
            {
            s->ck_pos.alloc(N); s->ck_vel.alloc(N); s->ck_last_pos.alloc(N); s->ck_accel.alloc(N); s->ck_image.alloc(N);
            }
        CUDA_OK(cudaMemcpyAsync(s->ck_pos.p, s->pos.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_OK(cudaMemcpyAsync(s->ck_vel.p, s->vel.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_OK(cudaMemcpyAsync(s->ck_accel.p, s->accel.p, sizeof(f3) * N, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_OK(cudaMemcpyAsync(s->ck_image.p, s->image.p, sizeof(i3) * N, cudaMemcpyDeviceToDevice, s->stream));
        StepState& st = *s->h_state;
        memset(&st, 0, sizeof(st));
        st.timestep = s->timestep;
        st.last_updated = s->timestep;
        st.every = 0x7fffffffu; // no rebuild decision can fire
        st.ntypes = s->ntypes;
        st.steps_left = which == 5 ? iters : iters + 1; // the multi-step launch ends with the chunk's (cheaper) last step
        st.N = s->N;
        float nvt_saved[4] = {0.0f, 0.0f, 1.0f, 0.0f};
        if (s->method == 3)
            {
            // the thermostat takes part in the scratch steps: start from "factor ready for the first launch", restore after
            CUDA_OK(cudaStreamSynchronize(s->stream));
            CUDA_OK(cudaMemcpy(nvt_saved, s->d_nvt.p, sizeof(nvt_saved), cudaMemcpyDeviceToHost));
            memcpy(&st.ts_step_two, &nvt_saved[3], 4);
            CUDA_OK(cudaMemsetAsync(s->nvt_acc.p, 0, sizeof(unsigned long long) * s->nvt_acc.n, s->stream));
            }
        CUDA_OK(cudaMemcpyAsync(s->d_state.p, s->h_state, sizeof(StepState), cudaMemcpyHostToDevice, s->stream));
        FusedStepArgs fa;
        fused_args(s, fa);
        fa.gs.tab = nullptr; // scratch steps must not write into the neighbors' ghost regions nor take the step barrier
        if (fa.mg) fa.mg = 1;
        CUDA_OK(cudaStreamSynchronize(s->stream));
        cudaError_t multi = cudaSuccess;
        CUDA_OK(cudaEventRecord(e0, s->stream));
        if (which == 5)
            multi = (s->pair_kind == B200_PAIR_LJ && !s->tile_on && s->bond_kind < 0)
                        ? launch_md_steps_lj(pa, (const float*)s->d_params.p, fa, s->pos.p, s->pos_alt.p, s->d_state.p, 0, 0ull, 0,
                                             s->stream)
                        : cudaErrorNotSupported;
        else
            for (unsigned int k = 0; k < iters; k++) enqueue_md_step(s, fa, (int)(k & 1), 0ull, 0);
        CUDA_OK(cudaEventRecord(e1, s->stream));
        CUDA_OK(cudaMemcpyAsync(s->pos.p, s->ck_pos.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_OK(cudaMemcpyAsync(s->vel.p, s->ck_vel.p, sizeof(float4) * N, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_OK(cudaMemcpyAsync(s->accel.p, s->ck_accel.p, sizeof(f3) * N, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_OK(cudaMemcpyAsync(s->image.p, s->ck_image.p, sizeof(i3) * N, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_OK(cudaStreamSynchronize(s->stream));
        if (s->method == 3) CUDA_OK(cudaMemcpy(s->d_nvt.p, nvt_saved, sizeof(nvt_saved), cudaMemcpyHostToDevice));
        float ms2 = 0.0f;
        CUDA_OK(cudaEventElapsedTime(&ms2, e0, e1));
        cudaGetLastError();
        if (multi != cudaSuccess) throw std::runtime_error("time_kernel(5): the multi-step kernel does not apply to this system");
        *avg_ms = (double)ms2 / iters;
        return 0;
        }
    CUDA_OK(cudaStreamSynchronize(s->stream));
    CUDA_OK(cudaEventRecord(e0, s->stream));
    for (unsigned int k = 0; k < iters; k++)
        {
        if (s->tile_on && which == 0)
            {
            FusedStepArgs fa;
            fused_args(s, fa);
            TileStepArgs ta;
            tile_step_args(s, ta, 1);
            CUDA_OK(launch_md_step_tile_lj(pa, (const float*)s->d_params.p, fa, ta, s->pos.p, nullptr, 0, s->d_state.p, 0ull, 0,
                                           s->stream));
            }
        else if (s->tile_on)
            tile_enqueue_rebuild(s, nullptr, s->pos.p); // identity permutation: the particles are in cell order
        else if (which == 0)
            CUDA_OK(b200_compute_pair_forces(s->pair_kind, &pa, pair_params_ptr(s), s->stream));
        else
            {
            // 1: the whole rebuild; 3: the cell list alone; 4: the neighbor-list kernel (+ exclusion filter) alone
            if (which != 4) CUDA_OK(launch_cell_list(&ca, nullptr, s->stream, s->mg ? &s->d_state.p->N : nullptr));
            if (which != 3)
                {
                CUDA_OK(launch_nlist_binned(&na, nullptr, s->stream));
                enqueue_nlist_filter(s, nullptr);
                }
            }
        }
    CUDA_OK(cudaEventRecord(e1, s->stream));
    CUDA_OK(cudaStreamSynchronize(s->stream));
    float ms = 0.0f;
    CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
    *avg_ms = (double)ms / iters;
    SYS_CATCH(s)
    }

int b200_sys_get_thermo(b200_system* s, double* out5)
    {
    SYS_TRY(s)
    const unsigned int nblocks = std::max(1u, std::min(1024u, (s->N + 255u) / 256u));
    CUDA_OK(b200_compute_thermo(s->th_out.p, s->th_partial.p, nblocks, (const float*)s->vel.p,
                                (const float*)s->net_force.p, s->compute_virial ? s->net_virial.p : nullptr,
                                s->virial_pitch, nullptr, s->N, s->stream));
    double h[9];
    s->th_out.download(h, 9, s->stream);
    // IntegratorTwoStep::getNDOF: D*N - D (hoomd/md/IntegratorTwoStep.cc:270-283); ComputeThermo stores FP32 properties
    const double ndof = 3.0 * method_group_size(s) - 3.0; // DOF of the integrated particles within group all
    const float ke = (float)h[0], pe = (float)h[1];
    const double volume = (double)(s->box.L[0] * s->box.L[1] * s->box.L[2]);
    const double W = (1.0f / 3.0f) * h[2];
    out5[0] = ke;
    out5[1] = pe;
    out5[2] = (float)(2.0f / (float)ndof) * ke;
    out5[3] = s->compute_virial ? (float)((2.0 * h[0] / 3.0 + W) / (float)volume) : 0.0;
    out5[4] = ndof;
    SYS_CATCH(s)
    }

} // extern "C"
