// PotentialPair<EvaluatorPairLJ> for sm_100a (replaces hoomd/md/PotentialPairGPU.cuh:208-471 + AllDriverPotentialPairGPU.cu:30-35).
// TPP lanes cooperate on one particle: they stride its CSR neighbor range (coalesced 4 B index loads), gather pos_j
// (16 B, L2-resident because particles are SFC-sorted), evaluate the pair on the FP32 FMA pipe and shuffle-reduce
// force / energy / virial; lane 0 of the group writes one float4 (+6 floats).  The evaluator is a template parameter,
// the same seam the reference uses (hoomd/md/PotentialPair.h:82-169), so other EvaluatorPair* can be plugged in.
#include "common.cuh"

namespace b200
{
// EvaluatorPairLJ::evalForceAndEnergy (hoomd/md/EvaluatorPairLJ.h:137-158)
struct EvaluatorPairLJ
    {
    typedef float2 param_type;
    __device__ __forceinline__ static bool eval(float rsq, float rcutsq, const float2& p, bool energy_shift,
                                                float& force_divr, float& pair_eng)
        {
        const float lj1 = p.x, lj2 = p.y;
        if (rsq < rcutsq && lj1 != 0.0f)
            {
            const float r2inv = __fdiv_rn(1.0f, rsq);
            const float r6inv = r2inv * r2inv * r2inv;
            force_divr = r2inv * r6inv * (12.0f * lj1 * r6inv - 6.0f * lj2);
            pair_eng = r6inv * (lj1 * r6inv - lj2);
            if (energy_shift)
                {
                const float rcut2inv = __fdiv_rn(1.0f, rcutsq);
                const float rcut6inv = rcut2inv * rcut2inv * rcut2inv;
                pair_eng -= rcut6inv * (lj1 * rcut6inv - lj2);
                }
            return true;
            }
        return false;
        }
    };

template<int TPP> __device__ __forceinline__ float group_reduce(float v)
    {
#pragma unroll
    for (int o = TPP / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
    }

template<class Evaluator, int TPP, int SHIFT_MODE, bool VIRIAL, bool MULTI_TYPE>
__global__ void __launch_bounds__(256) k_pair_forces(float4* __restrict__ force, float* __restrict__ virial,
                                                     unsigned int virial_pitch, unsigned int N,
                                                     const float4* __restrict__ pos, Box box,
                                                     const unsigned int* __restrict__ n_neigh,
                                                     const unsigned int* __restrict__ nlist,
                                                     const unsigned int* __restrict__ head_list,
                                                     const typename Evaluator::param_type* __restrict__ params,
                                                     const float* __restrict__ rcutsq_tab,
                                                     const float* __restrict__ ronsq_tab, unsigned int ntypes)
    {
    typedef typename Evaluator::param_type param_type;
    extern __shared__ unsigned char smem_raw[];
    param_type* s_params = reinterpret_cast<param_type*>(smem_raw);
    float* s_rcutsq = reinterpret_cast<float*>(s_params + (MULTI_TYPE ? ntypes * ntypes : 0));
    float* s_ronsq = s_rcutsq + (MULTI_TYPE ? ntypes * ntypes : 0);
    if (MULTI_TYPE)
        {
        for (unsigned int k = threadIdx.x; k < ntypes * ntypes; k += blockDim.x)
            {
            s_params[k] = params[k];
            s_rcutsq[k] = rcutsq_tab[k];
            s_ronsq[k] = (SHIFT_MODE == B200_SHIFT_XPLOR) ? ronsq_tab[k] : 0.0f;
            }
        __syncthreads();
        }
    const unsigned int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int i = gtid / TPP;
    const unsigned int sub = gtid % TPP;
    // no early exit: every lane of the warp takes part in the full-mask shuffles below (tail lanes idle with size 0)
    const bool active = i < N;
    const float4 pi4 = active ? pos[i] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    const unsigned int typei = (unsigned int)__float_as_int(pi4.w);
    const unsigned int size = active ? n_neigh[i] : 0u;
    const unsigned int head = active ? head_list[i] : 0u;

    param_type p0;
    float rcutsq0 = 0.0f, ronsq0 = 0.0f;
    if (!MULTI_TYPE)
        {
        p0 = params[0];
        rcutsq0 = rcutsq_tab[0];
        if (SHIFT_MODE == B200_SHIFT_XPLOR) ronsq0 = ronsq_tab[0];
        }

    float fx = 0.0f, fy = 0.0f, fz = 0.0f, pe = 0.0f;
    float vxx = 0.0f, vxy = 0.0f, vxz = 0.0f, vyy = 0.0f, vyz = 0.0f, vzz = 0.0f;

    // software-pipelined walk: the index of the next neighbor is in flight while this one is evaluated
    unsigned int k = sub;
    unsigned int next_j = (k < size) ? nlist[head + k] : 0u;
    for (; k < size; k += TPP)
        {
        const unsigned int j = next_j;
        if (k + TPP < size) next_j = nlist[head + k + TPP];
        const float4 pj4 = pos[j];
        float3 dx = make_float3(pi4.x - pj4.x, pi4.y - pj4.y, pi4.z - pj4.z);
        dx = min_image_fast(dx, box);
        const float rsq = dx.x * dx.x + dx.y * dx.y + dx.z * dx.z;

        param_type prm;
        float rcutsq, ronsq = 0.0f;
        if (MULTI_TYPE)
            {
            const unsigned int tp = typei * ntypes + (unsigned int)__float_as_int(pj4.w);
            prm = s_params[tp];
            rcutsq = s_rcutsq[tp];
            if (SHIFT_MODE == B200_SHIFT_XPLOR) ronsq = s_ronsq[tp];
            }
        else
            {
            prm = p0;
            rcutsq = rcutsq0;
            ronsq = ronsq0;
            }
        // hoomd/md/PotentialPair.h:413-424
        bool energy_shift = false;
        if (SHIFT_MODE == B200_SHIFT_SHIFT) energy_shift = true;
        else if (SHIFT_MODE == B200_SHIFT_XPLOR) energy_shift = ronsq > rcutsq;

        float force_divr = 0.0f, pair_eng = 0.0f;
        if (Evaluator::eval(rsq, rcutsq, prm, energy_shift, force_divr, pair_eng))
            {
            if (SHIFT_MODE == B200_SHIFT_XPLOR)
                {
                if (rsq >= ronsq && rsq < rcutsq) // hoomd/md/PotentialPair.h:441-465
                    {
                    const float old_pair_eng = pair_eng, old_force_divr = force_divr;
                    const float d = rcutsq - ronsq;
                    const float xplor_denom_inv = __fdiv_rn(1.0f, d * d * d);
                    const float rsq_minus_r_cut_sq = rsq - rcutsq;
                    const float s = rsq_minus_r_cut_sq * rsq_minus_r_cut_sq * (rcutsq + 2.0f * rsq - 3.0f * ronsq)
                                    * xplor_denom_inv;
                    const float ds_dr_divr = 12.0f * (rsq - ronsq) * rsq_minus_r_cut_sq * xplor_denom_inv;
                    pair_eng = old_pair_eng * s;
                    force_divr = s * old_force_divr - ds_dr_divr * old_pair_eng;
                    }
                }
            fx += dx.x * force_divr;
            fy += dx.y * force_divr;
            fz += dx.z * force_divr;
            pe += pair_eng * 0.5f;
            if (VIRIAL)
                {
                const float force_div2r = force_divr * 0.5f;
                vxx += force_div2r * dx.x * dx.x;
                vxy += force_div2r * dx.x * dx.y;
                vxz += force_div2r * dx.x * dx.z;
                vyy += force_div2r * dx.y * dx.y;
                vyz += force_div2r * dx.y * dx.z;
                vzz += force_div2r * dx.z * dx.z;
                }
            }
        }

    fx = group_reduce<TPP>(fx);
    fy = group_reduce<TPP>(fy);
    fz = group_reduce<TPP>(fz);
    pe = group_reduce<TPP>(pe);
    if (VIRIAL)
        {
        vxx = group_reduce<TPP>(vxx);
        vxy = group_reduce<TPP>(vxy);
        vxz = group_reduce<TPP>(vxz);
        vyy = group_reduce<TPP>(vyy);
        vyz = group_reduce<TPP>(vyz);
        vzz = group_reduce<TPP>(vzz);
        }
    if (sub == 0 && active)
        {
        force[i] = make_float4(fx, fy, fz, pe);
        if (VIRIAL)
            {
            virial[0 * virial_pitch + i] = vxx;
            virial[1 * virial_pitch + i] = vxy;
            virial[2 * virial_pitch + i] = vxz;
            virial[3 * virial_pitch + i] = vyy;
            virial[4 * virial_pitch + i] = vyz;
            virial[5 * virial_pitch + i] = vzz;
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
            a.d_head_list, params, a.d_rcutsq, a.d_ronsq, a.ntypes);
        }
    else
        k_pair_forces<Evaluator, TPP, SHIFT_MODE, VIRIAL, false><<<grid, bs, 0, stream>>>(
            (float4*)a.d_force, a.d_virial, a.virial_pitch, a.N, (const float4*)a.d_pos, box, a.d_n_neigh, a.d_nlist,
            a.d_head_list, params, a.d_rcutsq, a.d_ronsq, a.ntypes);
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

// default threads-per-particle: enough groups to fill 148 SMs x 2048 threads at small N, 1-2 lanes at large N where
// the kernel is bound by the nlist stream rather than by latency
unsigned int pick_pair_tpp(unsigned int N)
    {
    if (N <= 100000u) return 8;
    if (N <= 400000u) return 4;
    return 2;
    }

template<class Evaluator>
static cudaError_t launch_pair(const b200_pair_args& a, const typename Evaluator::param_type* params, cudaStream_t stream)
    {
    if (a.N == 0) return cudaSuccess;
    if (!box_is_orthorhombic(a.box)) return cudaErrorNotSupported;
    if (a.compute_virial && !a.d_virial) return cudaErrorInvalidValue;
    unsigned int tpp = a.threads_per_particle ? a.threads_per_particle : pick_pair_tpp(a.N);
    switch (tpp)
        {
        case 1: return launch_tpp<Evaluator, 1>(a, params, stream);
        case 2: return launch_tpp<Evaluator, 2>(a, params, stream);
        case 4: return launch_tpp<Evaluator, 4>(a, params, stream);
        case 8: return launch_tpp<Evaluator, 8>(a, params, stream);
        case 16: return launch_tpp<Evaluator, 16>(a, params, stream);
        case 32: return launch_tpp<Evaluator, 32>(a, params, stream);
        default: return cudaErrorInvalidValue;
        }
    }
} // namespace b200

extern "C" int b200_compute_lj_forces(const b200_pair_args* args, const float* d_params, b200_stream_t stream)
    {
    return (int)b200::launch_pair<b200::EvaluatorPairLJ>(*args, (const float2*)d_params, (cudaStream_t)stream);
    }
