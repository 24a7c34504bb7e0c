// Kernel template and launch ladder of PotentialPair<evaluator> (see pair_lj.cu for the description); included by one
// translation unit per evaluator so that the instantiations build in parallel.
#pragma once
#include "common.cuh"
#include "pair_body.cuh"
#include "kernels.h"

namespace b200
{
template<class Evaluator, int TPP, int SHIFT_MODE, bool VIRIAL, bool MULTI_TYPE>
__global__ void __launch_bounds__(256) k_pair_forces(float4* __restrict__ force, float* __restrict__ virial,
                                                     unsigned int virial_pitch, unsigned int N,
                                                     const float4* __restrict__ pos, Box box,
                                                     const unsigned int* __restrict__ n_neigh,
                                                     const unsigned int* __restrict__ nlist,
                                                     const unsigned int* __restrict__ head_list,
                                                     const typename Evaluator::param_type* __restrict__ params,
                                                     const float* __restrict__ rcutsq_tab,
                                                     const float* __restrict__ ronsq_tab, unsigned int ntypes,
                                                     float wrap_skip_margin)
    {
    extern __shared__ unsigned char smem_raw[];
    PairTables<Evaluator> T;
    pair_tables_load<Evaluator, SHIFT_MODE, MULTI_TYPE>(T, smem_raw, params, rcutsq_tab, ronsq_tab, ntypes);
    const unsigned int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int i = gtid / TPP;
    const unsigned int sub = gtid % TPP;
    // no early exit: every lane of the warp takes part in the full-mask shuffles (tail lanes idle with size 0)
    const bool active = i < N;
    const float4 pi4 = active ? pos[i] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    const unsigned int size = active ? n_neigh[i] : 0u;
    const unsigned int head = active ? head_list[i] : 0u;
    const PairAccum r = pair_accumulate<Evaluator, TPP, SHIFT_MODE, VIRIAL, MULTI_TYPE>(
        pi4, size, head, sub, pos, box, nlist, T, ntypes, !active || box_interior(pi4, box, wrap_skip_margin));
    if (sub == 0 && active)
        {
        force[i] = make_float4(r.fx, r.fy, r.fz, r.pe);
        if (VIRIAL)
            {
            virial[0 * virial_pitch + i] = r.vxx;
            virial[1 * virial_pitch + i] = r.vxy;
            virial[2 * virial_pitch + i] = r.vxz;
            virial[3 * virial_pitch + i] = r.vyy;
            virial[4 * virial_pitch + i] = r.vyz;
            virial[5 * virial_pitch + i] = r.vzz;
            }
        }
    }

template<class Evaluator, int TPP, int SHIFT_MODE, bool VIRIAL>
static cudaError_t launch_mt(const b200_pair_args& a, const typename Evaluator::param_type* params, cudaStream_t stream)
    {
    unsigned int bs = a.block_size ? a.block_size : 128;
    if (bs % 32 || bs > 256) return cudaErrorInvalidValue;
    const unsigned int per_block = bs / TPP;
    const unsigned int grid = (a.N + per_block - 1) / per_block;
    Box box = make_box(a.box);
    if (a.ntypes > 1)
        {
        size_t smem = (size_t)a.ntypes * a.ntypes * (sizeof(typename Evaluator::param_type) + 2 * sizeof(float));
        if (smem > 48 * 1024) return cudaErrorInvalidValue;
        k_pair_forces<Evaluator, TPP, SHIFT_MODE, VIRIAL, true><<<grid, bs, smem, stream>>>(
            (float4*)a.d_force, a.d_virial, a.virial_pitch, a.N, (const float4*)a.d_pos, box, a.d_n_neigh, a.d_nlist,
            a.d_head_list, params, a.d_rcutsq, a.d_ronsq, a.ntypes, a.wrap_skip_margin);
        }
    else
        k_pair_forces<Evaluator, TPP, SHIFT_MODE, VIRIAL, false><<<grid, bs, 0, stream>>>(
            (float4*)a.d_force, a.d_virial, a.virial_pitch, a.N, (const float4*)a.d_pos, box, a.d_n_neigh, a.d_nlist,
            a.d_head_list, params, a.d_rcutsq, a.d_ronsq, a.ntypes, a.wrap_skip_margin);
    return cudaGetLastError();
    }

template<class Evaluator, int TPP>
static cudaError_t launch_tpp(const b200_pair_args& a, const typename Evaluator::param_type* params, cudaStream_t stream)
    {
    const bool v = a.compute_virial != 0;
    switch (a.shift_mode)
        {
        case B200_SHIFT_NONE:
            return v ? launch_mt<Evaluator, TPP, B200_SHIFT_NONE, true>(a, params, stream)
                     : launch_mt<Evaluator, TPP, B200_SHIFT_NONE, false>(a, params, stream);
        case B200_SHIFT_SHIFT:
            return v ? launch_mt<Evaluator, TPP, B200_SHIFT_SHIFT, true>(a, params, stream)
                     : launch_mt<Evaluator, TPP, B200_SHIFT_SHIFT, false>(a, params, stream);
        case B200_SHIFT_XPLOR:
            return v ? launch_mt<Evaluator, TPP, B200_SHIFT_XPLOR, true>(a, params, stream)
                     : launch_mt<Evaluator, TPP, B200_SHIFT_XPLOR, false>(a, params, stream);
        default:
            return cudaErrorInvalidValue;
        }
    }

template<class Evaluator, bool ALL_TPP>
static cudaError_t launch_pair(const b200_pair_args& a, const typename Evaluator::param_type* params, cudaStream_t stream)
    {
    if (a.N == 0) return cudaSuccess;
    if (!box_is_orthorhombic(a.box)) return cudaErrorNotSupported;
    if (a.compute_virial && !a.d_virial) return cudaErrorInvalidValue;
    unsigned int tpp = a.threads_per_particle ? a.threads_per_particle : pick_pair_tpp(a.N);
    if (!ALL_TPP)
        {
        // evaluators beyond LJ are instantiated for 2 and 8 lanes per particle only (build time): round to the nearer
        if (tpp != 1 && tpp != 2 && tpp != 4 && tpp != 8 && tpp != 16 && tpp != 32) return cudaErrorInvalidValue;
        return tpp <= 2 ? launch_tpp<Evaluator, 2>(a, params, stream) : launch_tpp<Evaluator, 8>(a, params, stream);
        }
    switch (tpp)
        {
        case 1: return launch_tpp<Evaluator, ALL_TPP ? 1 : 2>(a, params, stream);
        case 2: return launch_tpp<Evaluator, 2>(a, params, stream);
        case 4: return launch_tpp<Evaluator, ALL_TPP ? 4 : 2>(a, params, stream);
        case 8: return launch_tpp<Evaluator, 8>(a, params, stream);
        case 16: return launch_tpp<Evaluator, ALL_TPP ? 16 : 8>(a, params, stream);
        case 32: return launch_tpp<Evaluator, ALL_TPP ? 32 : 8>(a, params, stream);
        default: return cudaErrorInvalidValue;
        }
    }
} // namespace b200
