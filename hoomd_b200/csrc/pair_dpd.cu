// PotentialPairDPDThermo<EvaluatorPairDPDThermo> -- md.pair.dpd, the pairwise DPD thermostat: conservative force
// A (1 - r/r_cut), drag -gamma w^2 (r^ . v_ij) and a random force sqrt(6 gamma kT / dt)-scaled per PAIR, drawn from a Saru
// stream keyed by (min tag, max tag, seed + timestep) so that both particles of a pair see the same number
// (hoomd/md/PotentialPairDPDThermo.h:116-270, hoomd/md/EvaluatorPairDPDThermo.h:161-230).  Unlike every other pair force it
// reads VELOCITIES and TAGS of the neighbors; energy and virial are the conservative part only.
//
// Lanes per particle stride the CSR row as in pair_body.cuh; the per-pair arithmetic follows the reference's CPU code
// operation by operation (1/sqrt as IEEE sqrt + division, separate roundings), so only the summation order differs.
#include "common.cuh"
#include "kernels.h"
#include "../../include/hoomd_b200.h"

namespace b200
{
constexpr int DPD_BLOCK = 128;

template<int TPP, bool VIRIAL, bool MULTI_TYPE>
__global__ void __launch_bounds__(DPD_BLOCK) k_pair_dpd_thermo(float4* __restrict__ force, float* __restrict__ virial,
                                                                unsigned int virial_pitch, unsigned int N,
                                                                const float4* __restrict__ pos,
                                                                const float4* __restrict__ vel,
                                                                const unsigned int* __restrict__ tag, Box box,
                                                                const unsigned int* __restrict__ n_neigh,
                                                                const unsigned int* __restrict__ nlist,
                                                                const unsigned int* __restrict__ head_list,
                                                                const float2* __restrict__ params,
                                                                const float* __restrict__ rcutsq_tab, unsigned int ntypes,
                                                                unsigned int seed, unsigned int timestep, float T,
                                                                float deltaT)
    {
    const unsigned int i = (blockIdx.x * DPD_BLOCK + threadIdx.x) / TPP;
    const unsigned int sub = threadIdx.x % TPP;
    const bool active = i < N;
    float fx = 0.0f, fy = 0.0f, fz = 0.0f, pe = 0.0f;
    float vir[6] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    if (active)
        {
        const float4 pi4 = pos[i];
        const float4 vi4 = vel[i];
        const unsigned int typei = (unsigned int)__float_as_int(pi4.w);
        const unsigned int tagi = tag[i];
        const unsigned int size = n_neigh[i];
        const unsigned int* row = nlist + head_list[i];
        for (unsigned int k = sub; k < size; k += TPP)
            {
            const unsigned int j = row[k];
            const float4 pj4 = pos[j];
            const float4 vj4 = vel[j];
            float3 dx = make_float3(__fsub_rn(pi4.x, pj4.x), __fsub_rn(pi4.y, pj4.y), __fsub_rn(pi4.z, pj4.z));
            const float3 dv = make_float3(__fsub_rn(vi4.x, vj4.x), __fsub_rn(vi4.y, vj4.y), __fsub_rn(vi4.z, vj4.z));
            dx = min_image_exact(dx, box);
            const float rsq = dot_exact(dx);
            const float rdotv = __fadd_rn(__fadd_rn(__fmul_rn(dx.x, dv.x), __fmul_rn(dx.y, dv.y)), __fmul_rn(dx.z, dv.z));
            const unsigned int tp = MULTI_TYPE ? typei * ntypes + (unsigned int)__float_as_int(pj4.w) : 0u;
            const float2 prm = params[tp];
            const float rcutsq = rcutsq_tab[tp];
            if (rsq < rcutsq)
                {
                const float a = prm.x, gamma = prm.y;
                const float rinv = __fdiv_rn(1.0f, __fsqrt_rn(rsq));
                const float r = __fdiv_rn(1.0f, rinv);
                const float rcutinv = __fdiv_rn(1.0f, __fsqrt_rn(rcutsq));
                const float rcut = __fdiv_rn(1.0f, rcutinv);
                const unsigned int tagj = tag[j];
                Saru rng(min(tagi, tagj), max(tagi, tagj), seed + timestep);
                const float alpha = rng.s(-1.0f, 1.0f);
                const float w = __fsub_rn(rinv, rcutinv);
                const float cons = __fmul_rn(a, w);
                float force_divr = cons;
                force_divr = __fsub_rn(force_divr, __fmul_rn(__fmul_rn(__fmul_rn(gamma, rdotv), w), w));
                const float amp = __fdiv_rn(1.0f, __fsqrt_rn(__fdiv_rn(deltaT, __fmul_rn(__fmul_rn(T, gamma), 6.0f))));
                force_divr = __fadd_rn(force_divr, __fmul_rn(__fmul_rn(amp, w), alpha));
                const float pair_eng = __fsub_rn(__fmul_rn(a, __fsub_rn(rcut, r)),
                                                 __fmul_rn(__fmul_rn(__fmul_rn(0.5f, a), rcutinv), __fsub_rn(rcutsq, rsq)));
                fx = __fmaf_rn(dx.x, force_divr, fx);
                fy = __fmaf_rn(dx.y, force_divr, fy);
                fz = __fmaf_rn(dx.z, force_divr, fz);
                pe = __fmaf_rn(pair_eng, 0.5f, pe);
                if (VIRIAL)
                    {
                    // conservative part only (PotentialPairDPDThermo.h:212-219): 0.5 * dx_a * dx_b * force_divr_cons
                    vir[0] = __fadd_rn(vir[0], __fmul_rn(__fmul_rn(__fmul_rn(0.5f, dx.x), dx.x), cons));
                    vir[1] = __fadd_rn(vir[1], __fmul_rn(__fmul_rn(__fmul_rn(0.5f, dx.x), dx.y), cons));
                    vir[2] = __fadd_rn(vir[2], __fmul_rn(__fmul_rn(__fmul_rn(0.5f, dx.x), dx.z), cons));
                    vir[3] = __fadd_rn(vir[3], __fmul_rn(__fmul_rn(__fmul_rn(0.5f, dx.y), dx.y), cons));
                    vir[4] = __fadd_rn(vir[4], __fmul_rn(__fmul_rn(__fmul_rn(0.5f, dx.y), dx.z), cons));
                    vir[5] = __fadd_rn(vir[5], __fmul_rn(__fmul_rn(__fmul_rn(0.5f, dx.z), dx.z), cons));
                    }
                }
            }
        }
    // reduce over the TPP lanes of the particle (all lanes of the warp take part)
#pragma unroll
    for (int o = TPP / 2; o > 0; o >>= 1)
        {
        fx += __shfl_xor_sync(0xffffffffu, fx, o);
        fy += __shfl_xor_sync(0xffffffffu, fy, o);
        fz += __shfl_xor_sync(0xffffffffu, fz, o);
        pe += __shfl_xor_sync(0xffffffffu, pe, o);
        if (VIRIAL)
            {
#pragma unroll
            for (int c = 0; c < 6; c++) vir[c] += __shfl_xor_sync(0xffffffffu, vir[c], o);
            }
        }
    if (active && sub == 0)
        {
        force[i] = make_float4(fx, fy, fz, pe);
        if (VIRIAL)
            {
#pragma unroll
            for (int c = 0; c < 6; c++) virial[(size_t)c * virial_pitch + i] = vir[c];
            }
        }
    }

template<int TPP>
static cudaError_t launch_dpd(const b200_pair_args& a, const float2* params, const b200_dpd_thermo_args& t,
                              cudaStream_t stream)
    {
    const unsigned int grid = (a.N * TPP + DPD_BLOCK - 1) / DPD_BLOCK;
    const Box box = make_box(a.box);
    const bool v = a.compute_virial != 0, mt = a.ntypes > 1;
#define B200_DPD_LAUNCH(V, M)                                                                                                 \
    k_pair_dpd_thermo<TPP, V, M><<<grid, DPD_BLOCK, 0, stream>>>(                                                            \
        (float4*)a.d_force, a.d_virial, a.virial_pitch, a.N, (const float4*)a.d_pos, (const float4*)t.d_vel, t.d_tag, box,     \
        a.d_n_neigh, a.d_nlist, a.d_head_list, params, a.d_rcutsq, a.ntypes, t.seed, t.timestep, t.T, t.deltaT)
    if (v && mt) B200_DPD_LAUNCH(true, true);
    else if (v) B200_DPD_LAUNCH(true, false);
    else if (mt) B200_DPD_LAUNCH(false, true);
    else B200_DPD_LAUNCH(false, false);
#undef B200_DPD_LAUNCH
    return cudaGetLastError();
    }
} // namespace b200

extern "C" int b200_compute_dpd_thermo_forces(const b200_pair_args* args, const float* d_params,
                                              const b200_dpd_thermo_args* thermo, b200_stream_t stream)
    {
    using namespace b200;
    if (!args || !thermo || !thermo->d_vel || !thermo->d_tag || !d_params) return (int)cudaErrorInvalidValue;
    if (args->N == 0) return 0;
    if (!box_is_orthorhombic(args->box)) return (int)cudaErrorNotSupported;
    if (!(thermo->T > 0.0f) || !(thermo->deltaT > 0.0f)) return (int)cudaErrorInvalidValue;
    const unsigned int tpp = args->threads_per_particle ? args->threads_per_particle : 4u;
    switch (tpp)
        {
        case 1: return (int)launch_dpd<1>(*args, (const float2*)d_params, *thermo, (cudaStream_t)stream);
        case 2: return (int)launch_dpd<2>(*args, (const float2*)d_params, *thermo, (cudaStream_t)stream);
        case 4: return (int)launch_dpd<4>(*args, (const float2*)d_params, *thermo, (cudaStream_t)stream);
        case 8: return (int)launch_dpd<8>(*args, (const float2*)d_params, *thermo, (cudaStream_t)stream);
        default: return (int)cudaErrorInvalidValue;
        }
    }
