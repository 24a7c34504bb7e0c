// Kernels of the device-driven step loop (engine only; group = all particles).
//   k_fused_step_one : IntegrationMethodTwoStep::integrateStepOne + NeighborList::distanceCheck in one pass over the
//                      particles -- the prologue of a chunk (the check re-uses the freshly written position: +16 B/particle
//                      instead of a second 32 B/particle kernel, hoomd/md/NeighborListGPU.cu:32-82)
//   k_decide         : NeighborList::needsUpdating (hoomd/md/NeighborList.cc:1080-1163) on the device, 1 thread
//   k_md_step        : the whole MD step in one launch (force + step two + next step one + check + decision)
//   (Nmax overflow after a device-side rebuild is flagged by the build kernel itself: OverflowHook, kernels.h)
// Per-particle arithmetic is the shared body of integrate_body.cuh / pair_body.cuh => bit-identical to the driver kernels.
#include "common.cuh"
#include "md_step_kernel.cuh"
#include "md_steps_persist.cuh"

namespace b200
{
__global__ void __launch_bounds__(256) k_fused_step_one(float4* __restrict__ pos, float4* __restrict__ vel,
                                                        const float3s* __restrict__ accel, int3* __restrict__ image,
                                                        const float4* __restrict__ last_pos,
                                                        const float* __restrict__ rcut_max, unsigned int N, Box box,
                                                        float dt, float r_buff, bool limit, float limit_val,
                                                        bool zero_force, StepState* __restrict__ st,
                                                        const unsigned int* __restrict__ d_N,
                                                        const NvtState* __restrict__ nvt, float3 lambda, float lambda_min,
                                                        const unsigned char* __restrict__ group_flag,
                                                        const unsigned int* __restrict__ tag)
    {
    unsigned int j = blockIdx.x * blockDim.x + threadIdx.x;
    bool moved = false;
    if (d_N) N = *d_N;
    if (j < N && !(group_flag && !group_flag[tag[j]])) // members of the method's group only
        {
        float4 p = pos[j];
        float4 v = vel[j];
        float3s a = accel[j];
        if (zero_force) a.x = a.y = a.z = 0.0f;
        int3 img = image[j];
        if (nvt) // TwoStepNVTMTK::integrateStepOne; the caller advances the thermostat afterwards
            nvt_step_one_body(p, v, a, img, box, dt, nvt->exp_fac);
        else
            nve_step_one_body(p, v, a, img, box, dt, limit, limit_val);
        pos[j] = p;
        vel[j] = v;
        image[j] = img;
        // NeighborList::distanceCheck, lambda = L / L_at_last_build (1 for a fixed box): hoomd/md/NeighborList.cc:981-1016
        const float4 lp = last_pos[j];
        const float old_rmin = rcut_max[(unsigned int)__float_as_int(p.w)];
        const float delta_max = __fdiv_rn(__fsub_rn(__fmul_rn(__fadd_rn(old_rmin, r_buff), lambda_min), old_rmin), 2.0f);
        const float maxsq = (delta_max > 0.0f) ? __fmul_rn(delta_max, delta_max) : 0.0f;
        float3 dx = make_float3(__fsub_rn(p.x, __fmul_rn(lambda.x, lp.x)), __fsub_rn(p.y, __fmul_rn(lambda.y, lp.y)),
                                __fsub_rn(p.z, __fmul_rn(lambda.z, lp.z)));
        dx = min_image_exact(dx, box);
        moved = dot_exact(dx) >= maxsq;
        }
    if (__any_sync(0xffffffffu, moved) && (threadIdx.x & 31) == 0) st->moved = 1u; // benign same-value race
    }

__global__ void k_decide(StepState* st)
    {
    decide_body(st);
    }

// first node of every replay of the step graph: hand the pending rebuild decision to the first conditional node
__global__ void k_graph_pre(StepState* st, cudaGraphConditionalHandle handle)
    {
#if (CUDART_VERSION >= 12040)
    cudaGraphSetConditional(handle, (st->steps_left && st->need_rebuild && !st->abort) ? 1u : 0u);
#endif
    }

__global__ void __launch_bounds__(256) k_compute_accel(float3s* __restrict__ accel,
                                                       const float4* __restrict__ net_force,
                                                       const float4* __restrict__ vel, unsigned int N)
    {
    unsigned int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= N) return;
    const float minv = __fdiv_rn(1.0f, vel[j].w);
    const float4 f = net_force[j];
    float3s a;
    a.x = __fmul_rn(f.x, minv);
    a.y = __fmul_rn(f.y, minv);
    a.z = __fmul_rn(f.z, minv);
    accel[j] = a;
    }

cudaError_t launch_fused_step_one(const FusedStepArgs& a, StepState* st, cudaStream_t stream)
    {
    if (a.N == 0) return cudaSuccess;
    // Langevin step one == NVE step one without limit / zero_force (hoomd/md/TwoStepLangevin.cc:86-125)
    const bool nve = a.method == 1;
    k_fused_step_one<<<(a.N + 255) / 256, 256, 0, stream>>>(a.pos, a.vel, (const float3s*)a.accel, (int3*)a.image,
                                                            a.last_pos, a.rcut_max, a.N, make_box(a.box), a.dt,
                                                            a.r_buff, nve && a.limit, a.limit_val, nve && a.zero_force,
                                                            st, a.d_N,
                                                            a.method == 3 ? (const NvtState*)a.nvt_state : nullptr,
                                                            make_float3(a.lambda[0], a.lambda[1], a.lambda[2]), a.lambda_min,
                                                            a.group_flag, a.tag);
    return cudaGetLastError();
    }

unsigned int md_step_block() { return (unsigned int)MD_BLOCK; }

cudaError_t launch_decide(StepState* st, cudaStream_t stream)
    {
    k_decide<<<1, 1, 0, stream>>>(st);
    return cudaGetLastError();
    }

cudaError_t launch_graph_pre(StepState* st, unsigned long long cond_handle, cudaStream_t stream)
    {
    k_graph_pre<<<1, 1, 0, stream>>>(st, (cudaGraphConditionalHandle)cond_handle);
    return cudaGetLastError();
    }

cudaError_t launch_md_step_lj(const b200_pair_args& pa, const float* d_params, const FusedStepArgs& fa,
                              const float4* pos_in, float4* pos_out, int copy_pos_on_last, StepState* st,
                              unsigned long long cond_handle, int use_cond, cudaStream_t stream)
    {
    return launch_md_step_t<EvaluatorPairLJ, true>(pa, d_params, fa, pos_in, pos_out, copy_pos_on_last, st, cond_handle,
                                                   use_cond, stream);
    }

cudaError_t launch_md_steps_lj(const b200_pair_args& pa, const float* d_params, const FusedStepArgs& fa, const float4* pos_a,
                               float4* pos_b, StepState* st, int probe_only, unsigned long long next_handle, int use_cond,
                               cudaStream_t stream)
    {
    return launch_md_steps_t<EvaluatorPairLJ>(pa, d_params, fa, pos_a, pos_b, st, probe_only != 0, next_handle, use_cond,
                                              stream);
    }

// the same step with another pair evaluator (pair_ev.cu holds the instantiations)
cudaError_t launch_md_step_kind(int evaluator, const b200_pair_args& pa, const void* d_params, const FusedStepArgs& fa,
                                const float4* pos_in, float4* pos_out, int copy_pos_on_last, StepState* st,
                                unsigned long long cond_handle, int use_cond, cudaStream_t stream)
    {
    switch (evaluator)
        {
        case B200_PAIR_LJ:
            return launch_md_step_lj(pa, (const float*)d_params, fa, pos_in, pos_out, copy_pos_on_last, st, cond_handle,
                                     use_cond, stream);
        case B200_PAIR_GAUSS:
            return launch_md_step_gauss(pa, d_params, fa, pos_in, pos_out, copy_pos_on_last, st, cond_handle, use_cond, stream);
        case B200_PAIR_YUKAWA:
            return launch_md_step_yukawa(pa, d_params, fa, pos_in, pos_out, copy_pos_on_last, st, cond_handle, use_cond, stream);
        case B200_PAIR_MIE:
            return launch_md_step_mie(pa, d_params, fa, pos_in, pos_out, copy_pos_on_last, st, cond_handle, use_cond, stream);
        case B200_PAIR_FORCE_SHIFTED_LJ:
            return launch_md_step_fslj(pa, d_params, fa, pos_in, pos_out, copy_pos_on_last, st, cond_handle, use_cond, stream);
        case B200_PAIR_MORSE:
            return launch_md_step_morse(pa, d_params, fa, pos_in, pos_out, copy_pos_on_last, st, cond_handle, use_cond, stream);
        case B200_PAIR_DPD:
            return launch_md_step_dpd(pa, d_params, fa, pos_in, pos_out, copy_pos_on_last, st, cond_handle, use_cond, stream);
        default:
            return cudaErrorInvalidValue;
        }
    }

cudaError_t launch_compute_accel(float* accel, const float4* net_force, const float4* vel, unsigned int N,
                                 cudaStream_t stream)
    {
    if (N == 0) return cudaSuccess;
    k_compute_accel<<<(N + 255) / 256, 256, 0, stream>>>((float3s*)accel, net_force, vel, N);
    return cudaGetLastError();
    }
} // namespace b200

// out[i] = (xyz of a[i], w of b[i]) -- rebuilds the positions a neighbor list was built from (last_pos carries no type)
namespace b200
{
__global__ void __launch_bounds__(256) k_merge_xyz_w(float4* __restrict__ out, const float4* __restrict__ a,
                                                     const float4* __restrict__ b, unsigned int N)
    {
    unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const float4 x = a[i];
    out[i] = make_float4(x.x, x.y, x.z, b[i].w);
    }

cudaError_t launch_merge_xyz_w(float4* out, const float4* a, const float4* b, unsigned int N, cudaStream_t stream)
    {
    if (N == 0) return cudaSuccess;
    k_merge_xyz_w<<<(N + 255) / 256, 256, 0, stream>>>(out, a, b, N);
    return cudaGetLastError();
    }

// ordered compaction by one CTA (runs after a particle sort only): thread t owns a contiguous range of indices
__global__ void __launch_bounds__(1024) k_group_index(unsigned int* __restrict__ idx, unsigned int* __restrict__ count,
                                                      const unsigned char* __restrict__ group_flag,
                                                      const unsigned int* __restrict__ tag, unsigned int N)
    {
    __shared__ unsigned int s_warp[32];
    const unsigned int per = (N + 1023u) / 1024u;
    const unsigned int b = min(N, threadIdx.x * per), e = min(N, b + per);
    unsigned int cnt = 0;
    for (unsigned int k = b; k < e; k++) cnt += group_flag[tag[k]] ? 1u : 0u;
    unsigned int incl = cnt;
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
        {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((int)lane >= o) incl += t;
        }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0)
        {
        const unsigned int w = s_warp[lane];
        unsigned int wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
            {
            const unsigned int t = __shfl_up_sync(0xffffffffu, wi, o);
            if ((int)lane >= o) wi += t;
            }
        s_warp[lane] = wi - w;
        if (lane == 31) *count = wi;
        }
    __syncthreads();
    unsigned int o = s_warp[warp] + incl - cnt;
    for (unsigned int k = b; k < e; k++)
        if (group_flag[tag[k]]) idx[o++] = k;
    }

cudaError_t launch_group_index(unsigned int* idx, unsigned int* count, const unsigned char* group_flag,
                               const unsigned int* tag, unsigned int N, cudaStream_t stream)
    {
    k_group_index<<<1, 1024, 0, stream>>>(idx, count, group_flag, tag, N);
    return cudaGetLastError();
    }

__global__ void __launch_bounds__(256) k_gather_by_tag(float* __restrict__ out, const float* __restrict__ by_tag,
                                                       const unsigned int* __restrict__ tag, unsigned int N)
    {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) out[i] = by_tag[tag[i]];
    }

cudaError_t launch_gather_by_tag(float* out, const float* by_tag, const unsigned int* tag, unsigned int N,
                                 cudaStream_t stream)
    {
    if (N == 0) return cudaSuccess;
    k_gather_by_tag<<<(N + 255) / 256, 256, 0, stream>>>(out, by_tag, tag, N);
    return cudaGetLastError();
    }

__global__ void k_tally_fold(double* __restrict__ acc, const float* __restrict__ e, unsigned int ts)
    {
    acc[0] += (double)*e;
    acc[1u + (ts & 1u)] = (double)*e;
    }

cudaError_t launch_tally_fold(double* acc, const float* e, unsigned int ts, cudaStream_t stream)
    {
    k_tally_fold<<<1, 1, 0, stream>>>(acc, e, ts);
    return cudaGetLastError();
    }
} // namespace b200

#ifdef B200_PERSIST_TRACE
// debugging builds only (see md_steps_persist.cuh): copy the phase stamps out and reset the per-CTA step counters
extern "C" __attribute__((visibility("default"))) int b200_debug_persist_trace(unsigned long long* host, unsigned int* dims)
    {
    using namespace b200;
    dims[0] = PERSIST_TRACE_CTAS; dims[1] = PERSIST_TRACE_STEPS; dims[2] = PERSIST_TRACE_MARKS;
    cudaError_t e = cudaDeviceSynchronize();
    if (host && e == cudaSuccess)
        e = cudaMemcpyFromSymbol(host, g_persist_trace, sizeof(unsigned long long) * PERSIST_TRACE_CTAS * PERSIST_TRACE_STEPS
                                                             * PERSIST_TRACE_MARKS);
    static unsigned int zeros[PERSIST_TRACE_CTAS];
    if (e == cudaSuccess) e = cudaMemcpyToSymbol(g_persist_trace_step, zeros, sizeof(zeros));
    return (int)e;
    }
#endif

#ifdef B200_FAULT_INJECT
// fault-injection builds only (see md_steps_persist.cuh): the next step of k_md_steps loses the ticket of CTA `cta`
extern "C" __attribute__((visibility("default"))) int b200_debug_fault_lose_cta(unsigned int cta)
    {
    const unsigned int v = cta + 1u;
    return (int)cudaMemcpyToSymbol(b200::g_fault_lose_cta, &v, sizeof(v));
    }
#endif
