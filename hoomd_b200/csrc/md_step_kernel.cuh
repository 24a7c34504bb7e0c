// Kernel template and launch ladder of the single-kernel MD step (see fused_step.cu for the description); included by one
// translation unit per evaluator so that the instantiations build in parallel.
#pragma once
#include "common.cuh"
#include "md_step_body.cuh"
#include <cstdlib>

#include <type_traits>

namespace b200
{
// register bound for the plain variants (single type, no virial, no xplor, LJ): 10 CTAs of 128 threads per SM (48 registers,
// no spills), or OCC = 1: 12 CTAs (40 registers, 36 B of spills outside the neighbor loop).  Measured (profiles/r1r_*.json):
// at 1M particles the extra warps hide more gather latency (4257 -> 4388 TPS), at 64k the spills cost (31.4k -> 31.0k), so the
// launcher picks OCC by system size.  The multi-type / virial / xplor variants and the exp/pow evaluators keep their 56-62
// registers.
template<class Evaluator, int SHIFT_MODE, bool VIRIAL, bool MULTI_TYPE> struct MdStepPlain
    {
    static constexpr bool value = !MULTI_TYPE && !VIRIAL && SHIFT_MODE != 2 && std::is_same<Evaluator, EvaluatorPairLJ>::value;
    };

template<class Evaluator, int TPP, int SHIFT_MODE, bool VIRIAL, bool MULTI_TYPE, int OCC = 0, bool EXTRAS = true>
__global__ void __launch_bounds__(MD_BLOCK, !MdStepPlain<Evaluator, SHIFT_MODE, VIRIAL, MULTI_TYPE>::value
                                                ? 1 : (OCC ? 1536 / MD_BLOCK : 1280 / MD_BLOCK))
k_md_step(const __grid_constant__ MdStepArgs<Evaluator> a)
    {
    constexpr int PPB = MD_BLOCK / TPP; // particles per CTA
    extern __shared__ unsigned char smem_raw[];
    __shared__ float4 s_force[PPB];
    __shared__ float s_virial[VIRIAL ? 6 * PPB : 1];
    StepState* st = a.st;
    const unsigned int steps_left = st->steps_left;
    if (steps_left == 0 || st->abort)
        {
        // padding launch of an odd-length chunk / aborted chunk (grid-uniform).  After a collective abort (Nmax overflow
        // vote) nothing may run until the host has grown Nmax -- in particular not another rebuild bracket, which would
        // migrate from the other, older position buffer: withdraw the pending rebuild request.
        if (st->abort && blockIdx.x == 0 && threadIdx.x == 0) st->need_rebuild = 0u;
        return;
        }
    const bool last = steps_left == 1;
    const unsigned int ts2 = st->ts_step_two;
    const bool store_force = last || st->store_force;
    const unsigned int N = (EXTRAS && a.d_N) ? *a.d_N : a.N;
    // multi-GPU: chunks that touch the ghost layer are scheduled first (their ghost stores overlap the interior chunks)
    const unsigned int chunk = (EXTRAS && a.gs.cta_order) ? a.gs.cta_order[blockIdx.x] : blockIdx.x;
    __shared__ float s_exp_fac;
    md_step_nvt_begin(a, st, ts2); // NVT: thread 0 of block 0 advances the thermostat for this launch (no-op otherwise)

    // ---- phase 1: TPP lanes per particle walk its neighbor list (PotentialPair::computeForces) ----
    PairTables<Evaluator> T;
    pair_tables_load<Evaluator, SHIFT_MODE, MULTI_TYPE>(T, smem_raw, a.params, a.rcutsq, a.ronsq, a.ntypes);
        {
        const unsigned int slot = threadIdx.x / TPP;
        const unsigned int i = chunk * PPB + slot;
        const unsigned int sub = threadIdx.x % TPP;
        const bool active = i < N;
        if (a.prefetch && threadIdx.x < PPB)
            {
            // operands of the integration phase of particle chunk*PPB + threadIdx.x: in L1 by the time the walk is done
            const unsigned int i2 = chunk * PPB + threadIdx.x;
            if (i2 < N)
                {
                prefetch_l1(a.vel + i2);
                prefetch_l1(a.image + i2);
                prefetch_l1(a.last_pos + i2);
                }
            }
        const float4 pi4 = active ? a.pos_in[i] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        const unsigned int size = active ? a.n_neigh[i] : 0u;
        // one type: the head list is the exclusive scan of a constant, no load on the way to the first index chunk
        const unsigned int head = !active ? 0u : ((!MULTI_TYPE && a.nmax_uniform) ? i * a.nmax_uniform : a.head_list[i]);
        const PairAccum r = pair_accumulate<Evaluator, TPP, SHIFT_MODE, VIRIAL, MULTI_TYPE>(
            pi4, size, head, sub, a.pos_in, a.box, a.nlist, T, a.ntypes,
            !active || box_interior(pi4, a.box, a.wrap_skip_margin));
        if (sub == 0)
            {
            s_force[slot] = make_float4(r.fx, r.fy, r.fz, r.pe);
            if (VIRIAL)
                {
                s_virial[0 * PPB + slot] = r.vxx;
                s_virial[1 * PPB + slot] = r.vxy;
                s_virial[2 * PPB + slot] = r.vxz;
                s_virial[3 * PPB + slot] = r.vyy;
                s_virial[4 * PPB + slot] = r.vyz;
                s_virial[5 * PPB + slot] = r.vzz;
                }
            }
        }
    md_step_nvt_wait(a, st, ts2, &s_exp_fac);
    __syncthreads();

    // ---- phase 2: one lane per particle integrates (full warps, coalesced) instead of one lane in TPP ----
    bool moved = false, pushed = false;
    double mv2 = 0.0;
    const unsigned int i = chunk * PPB + threadIdx.x;
    if (threadIdx.x < PPB && i < N)
        {
        float vir[6] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
        if (VIRIAL)
            {
#pragma unroll
            for (int c = 0; c < 6; c++) vir[c] = s_virial[c * PPB + threadIdx.x];
            }
        moved = md_step_integrate<Evaluator, VIRIAL, EXTRAS>(a, i, s_force[threadIdx.x], a.pos_in[i], vir, ts2, last,
                                                             store_force, mv2, a.method == 3 ? s_exp_fac : 1.0f, &pushed);
        }
    md_step_finish<Evaluator, EXTRAS>(a, st, moved, steps_left, last, mv2, ts2, pushed);
    }

template<class Evaluator, int TPP, int SHIFT_MODE, bool VIRIAL>
static cudaError_t md_step_mt(const MdStepArgs<Evaluator>& a, cudaStream_t stream)
    {
    // L1 / shared split: the kernel needs <= 10 kB of shared memory; the neighbor-position gathers live off L1
    static const int carveout = getenv("HOOMD_B200_CARVEOUT") ? atoi(getenv("HOOMD_B200_CARVEOUT")) : -1;
    static bool configured = false;
    if (!configured && carveout >= 0)
        {
        cudaFuncSetAttribute(k_md_step<Evaluator, TPP, SHIFT_MODE, VIRIAL, true>,
                             cudaFuncAttributePreferredSharedMemoryCarveout, carveout);
        cudaFuncSetAttribute(k_md_step<Evaluator, TPP, SHIFT_MODE, VIRIAL, false>,
                             cudaFuncAttributePreferredSharedMemoryCarveout, carveout);
        configured = true;
        }
    const unsigned int bs = MD_BLOCK;
    const unsigned int per_block = bs / TPP;
    const unsigned int grid = (a.N + per_block - 1) / per_block;
    if (a.ntypes > 1)
        {
        size_t smem = (size_t)a.ntypes * a.ntypes * (sizeof(typename Evaluator::param_type) + 2 * sizeof(float));
        if (smem > 32 * 1024) return cudaErrorInvalidValue;
        k_md_step<Evaluator, TPP, SHIFT_MODE, VIRIAL, true><<<grid, bs, smem, stream>>>(a);
        }
    else
        {
        if constexpr (MdStepPlain<Evaluator, SHIFT_MODE, VIRIAL, false>::value)
            {
            // large systems are latency bound on the neighbor gathers: trade 8 registers for 2 more CTAs per SM
            static const char* occ_env = getenv("HOOMD_B200_MD_OCC");
            const bool high_occ = occ_env ? atoi(occ_env) != 0 : a.N >= 200000u;
            // group / dscale / tally / domain decomposition: the variants that carry those paths (a few registers more)
            const bool extras = a.group_flag || a.tally_acc || a.use_lambda || a.mg || a.gs.tab || a.gs.cta_order || a.d_N;
            if (high_occ)
                {
                if (extras)
                    k_md_step<Evaluator, TPP, SHIFT_MODE, VIRIAL, false, 1, true><<<grid, bs, 0, stream>>>(a);
                else
                    k_md_step<Evaluator, TPP, SHIFT_MODE, VIRIAL, false, 1, false><<<grid, bs, 0, stream>>>(a);
                return cudaGetLastError();
                }
            if (!extras)
                {
                k_md_step<Evaluator, TPP, SHIFT_MODE, VIRIAL, false, 0, false><<<grid, bs, 0, stream>>>(a);
                return cudaGetLastError();
                }
            }
        k_md_step<Evaluator, TPP, SHIFT_MODE, VIRIAL, false><<<grid, bs, 0, stream>>>(a);
        }
    return cudaGetLastError();
    }

template<class Evaluator, int TPP>
static cudaError_t md_step_tpp(const MdStepArgs<Evaluator>& a, unsigned int shift_mode, bool v, cudaStream_t stream)
    {
    switch (shift_mode)
        {
        case B200_SHIFT_NONE:
            return v ? md_step_mt<Evaluator, TPP, B200_SHIFT_NONE, true>(a, stream)
                     : md_step_mt<Evaluator, TPP, B200_SHIFT_NONE, false>(a, stream);
        case B200_SHIFT_SHIFT:
            return v ? md_step_mt<Evaluator, TPP, B200_SHIFT_SHIFT, true>(a, stream)
                     : md_step_mt<Evaluator, TPP, B200_SHIFT_SHIFT, false>(a, stream);
        case B200_SHIFT_XPLOR:
            return v ? md_step_mt<Evaluator, TPP, B200_SHIFT_XPLOR, true>(a, stream)
                     : md_step_mt<Evaluator, TPP, B200_SHIFT_XPLOR, false>(a, stream);
        default:
            return cudaErrorInvalidValue;
        }
    }

// the launch ladder for one evaluator; ALL_TPP = false instantiates 2 and 8 lanes per particle only (build time)
template<class Evaluator, bool ALL_TPP>
static cudaError_t launch_md_step_t(const b200_pair_args& pa, const void* d_params, const FusedStepArgs& fa,
                                    const float4* pos_in, float4* pos_out, int copy_pos_on_last, StepState* st,
                                    unsigned long long cond_handle, int use_cond, cudaStream_t stream)
    {
    if (fa.N == 0) return cudaSuccess;
    if (!box_is_orthorhombic(pa.box)) return cudaErrorNotSupported;
    MdStepArgs<Evaluator> a;
    fill_md_step_args(a, pa, (const float*)d_params, fa, pos_in, pos_out, copy_pos_on_last, st, cond_handle, use_cond);
    const bool v = pa.compute_virial != 0;
    const unsigned int tpp = pa.threads_per_particle ? pa.threads_per_particle : pick_pair_tpp(pa.N);
    if (!ALL_TPP)
        return tpp <= 2 ? md_step_tpp<Evaluator, 2>(a, pa.shift_mode, v, stream)
                        : md_step_tpp<Evaluator, 8>(a, pa.shift_mode, v, stream);
    switch (tpp)
        {
        case 1: return md_step_tpp<Evaluator, ALL_TPP ? 1 : 2>(a, pa.shift_mode, v, stream);
        case 2: return md_step_tpp<Evaluator, 2>(a, pa.shift_mode, v, stream);
        case 4: return md_step_tpp<Evaluator, ALL_TPP ? 4 : 2>(a, pa.shift_mode, v, stream);
        case 8: return md_step_tpp<Evaluator, 8>(a, pa.shift_mode, v, stream);
        case 16: return md_step_tpp<Evaluator, ALL_TPP ? 16 : 8>(a, pa.shift_mode, v, stream);
        case 32: return md_step_tpp<Evaluator, ALL_TPP ? 32 : 8>(a, pa.shift_mode, v, stream);
        default: return cudaErrorInvalidValue;
        }
    }
} // namespace b200
