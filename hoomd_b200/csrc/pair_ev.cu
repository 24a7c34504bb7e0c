// One translation unit per pair evaluator beyond LJ (compiled with -DB200_EV=<n>, see the Makefile): the stand-alone
// PotentialPair<evaluator> force kernel and the single-kernel MD step, instantiated through the same templates as LJ.
// Reference: hoomd/md/AllPairPotentials.h:45-79, hoomd/md/AllDriverPotentialPairGPU.cu.
#include "pair_kernel.cuh"
#include "md_step_kernel.cuh"

#if B200_EV == 1
#define EV EvaluatorPairGauss
#define NAME gauss
#elif B200_EV == 2
#define EV EvaluatorPairYukawa
#define NAME yukawa
#elif B200_EV == 3
#define EV EvaluatorPairMie
#define NAME mie
#elif B200_EV == 4
#define EV EvaluatorPairForceShiftedLJ
#define NAME fslj
#elif B200_EV == 5
#define EV EvaluatorPairMorse
#define NAME morse
#elif B200_EV == 6
#define EV EvaluatorPairDPD
#define NAME dpd
#else
#error "B200_EV must be 1..6"
#endif
#define CAT2(a, b) a##b
#define CAT(a, b) CAT2(a, b)

namespace b200
{
cudaError_t CAT(launch_pair_, NAME)(const b200_pair_args& a, const void* d_params, cudaStream_t stream)
    {
    return launch_pair<EV, false>(a, (const EV::param_type*)d_params, stream);
    }

cudaError_t CAT(launch_md_step_, NAME)(const b200_pair_args& pa, const void* d_params, const FusedStepArgs& fa,
                                       const float4* pos_in, float4* pos_out, int copy_pos_on_last, StepState* st,
                                       unsigned long long cond_handle, int use_cond, cudaStream_t stream)
    {
    return launch_md_step_t<EV, false>(pa, d_params, fa, pos_in, pos_out, copy_pos_on_last, st, cond_handle, use_cond,
                                       stream);
    }
} // namespace b200
