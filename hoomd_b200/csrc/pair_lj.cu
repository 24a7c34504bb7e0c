// PotentialPair<EvaluatorPairLJ> for sm_100a (replaces hoomd/md/PotentialPairGPU.cuh:208-471 + AllDriverPotentialPairGPU.cu:30-35).
// TPP lanes cooperate on one particle: they stride its CSR neighbor range (coalesced 4 B index loads), gather pos_j
// (16 B, L2-resident because particles are SFC-sorted), evaluate the pair on the FP32 FMA pipe and shuffle-reduce
// force / energy / virial; lane 0 of the group writes one float4 (+6 floats).  The evaluator is a template parameter,
// the same seam the reference uses (hoomd/md/PotentialPair.h:82-169), so other EvaluatorPair* can be plugged in.
#include "pair_kernel.cuh"

namespace b200
{
// default threads-per-particle (measured on B200, LJ liquid rho 0.2: 64k particles 29.4k / 27.3k / 25.2k steps/s at
// TPP 2 / 4 / 8; 1M particles 3.7k / 3.2k at TPP 2 / 4): 2 lanes once N fills the 148 SMs, more lanes for small systems
unsigned int pick_pair_tpp(unsigned int N)
    {
    if (N <= 12000u) return 16;
    if (N <= 24000u) return 8;
    if (N <= 48000u) return 4;
    return 2;
    }

} // namespace b200

extern "C" int b200_compute_lj_forces(const b200_pair_args* args, const float* d_params, b200_stream_t stream)
    {
    return (int)b200::launch_pair<b200::EvaluatorPairLJ, true>(*args, (const float2*)d_params, (cudaStream_t)stream);
    }

// PotentialPair<evaluator> for the other evaluators of AllPairPotentials.h (one translation unit each: pair_ev.cu)
extern "C" int b200_compute_pair_forces(int evaluator, const b200_pair_args* args, const void* d_params,
                                        b200_stream_t stream)
    {
    using namespace b200;
    if (!args || !d_params) return (int)cudaErrorInvalidValue;
    switch (evaluator)
        {
        case B200_PAIR_LJ: return b200_compute_lj_forces(args, (const float*)d_params, stream);
        case B200_PAIR_GAUSS: return (int)launch_pair_gauss(*args, d_params, (cudaStream_t)stream);
        case B200_PAIR_YUKAWA: return (int)launch_pair_yukawa(*args, d_params, (cudaStream_t)stream);
        case B200_PAIR_MIE: return (int)launch_pair_mie(*args, d_params, (cudaStream_t)stream);
        case B200_PAIR_FORCE_SHIFTED_LJ: return (int)launch_pair_fslj(*args, d_params, (cudaStream_t)stream);
        case B200_PAIR_MORSE: return (int)launch_pair_morse(*args, d_params, (cudaStream_t)stream);
        case B200_PAIR_DPD: return (int)launch_pair_dpd(*args, d_params, (cudaStream_t)stream);
        default: return (int)cudaErrorInvalidValue;
        }
    }
