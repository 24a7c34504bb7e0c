// Integrator kernels for sm_100a: net-force sum, TwoStepNVE step one/two, TwoStepLangevin step two, Saru probe.
// These are pure streaming kernels (HBM-bound, ~60-100 B per particle); one thread per group member,
// 16 B vector loads/stores for pos/vel/net_force.  Arithmetic follows the CPU oracle operation by operation
// (hoomd/md/TwoStepNVE.cc:77-122,255-300, hoomd/md/TwoStepLangevin.cc:86-125,240-330) with explicit *_rn intrinsics so
// nothing is contracted into FMAs: one half step is bit-identical to the reference.
#include "common.cuh"
#include "integrate_body.cuh"
#include "kernels.h"
#include <cstdio>
#include <mutex>

namespace b200
{

// ---- Integrator::computeNetForceGPU (hoomd/Integrator.cu:57-113) --------------------------------------------------
template<bool VIRIAL>
__global__ void __launch_bounds__(256) k_sum_net_force(float4* __restrict__ net_force, float* __restrict__ net_virial,
                                                       unsigned int net_virial_pitch, b200_force_list fl,
                                                       unsigned int n, int clear)
    {
    unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 f = clear ? make_float4(0.0f, 0.0f, 0.0f, 0.0f) : net_force[i];
    float v[6] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    if (VIRIAL && !clear)
        {
#pragma unroll
        for (int k = 0; k < 6; k++) v[k] = net_virial[k * net_virial_pitch + i];
        }
#pragma unroll
    for (int a = 0; a < 6; a++)
        {
        if (fl.f[a])
            {
            const float4 g = reinterpret_cast<const float4*>(fl.f[a])[i];
            f.x = __fadd_rn(f.x, g.x);
            f.y = __fadd_rn(f.y, g.y);
            f.z = __fadd_rn(f.z, g.z);
            f.w = __fadd_rn(f.w, g.w);
            if (VIRIAL && fl.v[a])
                {
#pragma unroll
                for (int k = 0; k < 6; k++) v[k] = __fadd_rn(v[k], fl.v[a][k * fl.vpitch[a] + i]);
                }
            }
        }
    net_force[i] = f;
    if (VIRIAL)
        {
#pragma unroll
        for (int k = 0; k < 6; k++) net_virial[k * net_virial_pitch + i] = v[k];
        }
    }

__global__ void __launch_bounds__(256) k_nve_step_one(float4* __restrict__ pos, float4* __restrict__ vel,
                                                      const float3s* __restrict__ accel, int3* __restrict__ image,
                                                      const unsigned int* __restrict__ group, unsigned int group_size,
                                                      Box box, float dt, bool limit, float limit_val, bool zero_force)
    {
    unsigned int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= group_size) return;
    const unsigned int j = group ? group[g] : g;
    float4 p = pos[j];
    float4 v = vel[j];
    float3s a = accel[j];
    if (zero_force) a.x = a.y = a.z = 0.0f;
    int3 img = image[j];
    nve_step_one_body(p, v, a, img, box, dt, limit, limit_val);
    pos[j] = p;
    vel[j] = v;
    image[j] = img;
    }

__global__ void __launch_bounds__(256) k_nve_step_two(float4* __restrict__ vel, float3s* __restrict__ accel,
                                                      const unsigned int* __restrict__ group, unsigned int group_size,
                                                      const float4* __restrict__ net_force, float dt, bool limit,
                                                      float limit_val, bool zero_force)
    {
    unsigned int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= group_size) return;
    const unsigned int j = group ? group[g] : g;
    float4 v = vel[j];
    float3s a = accel[j];
    nve_step_two_body(v, a, net_force[j], dt, limit, limit_val, zero_force);
    vel[j] = v;
    accel[j] = a;
    }

__global__ void __launch_bounds__(256) k_langevin_step_two(const float4* __restrict__ pos, float4* __restrict__ vel,
                                                           float3s* __restrict__ accel,
                                                           const float* __restrict__ diameter,
                                                           const unsigned int* __restrict__ tag,
                                                           const unsigned int* __restrict__ group,
                                                           unsigned int group_size,
                                                           const float4* __restrict__ net_force,
                                                           const float* __restrict__ gamma_tab, bool use_lambda,
                                                           float lambda, float T, unsigned int timestep,
                                                           unsigned int seed, float* __restrict__ sum_bdenergy,
                                                           bool noiseless_t, bool tally, float dt, unsigned int D)
    {
    unsigned int g = blockIdx.x * blockDim.x + threadIdx.x;
    float e = 0.0f;
    if (g < group_size)
        {
        const unsigned int j = group ? group[g] : g;
        float4 v = vel[j];
        float3s a;
        float gamma;
        if (use_lambda) gamma = __fmul_rn(lambda, diameter[j]);
        else gamma = gamma_tab[(unsigned int)__float_as_int(pos[j].w)];
        e = langevin_step_two_body(v, a, net_force[j], tag[j], timestep, seed, gamma, T, dt, noiseless_t, D);
        vel[j] = v;
        accel[j] = a;
        }
    if (tally)
        {
        // warp shuffle reduce, one atomic per warp (the reference block-reduces then sums partials,
        // hoomd/md/TwoStepLangevinGPU.cu:150-240; the tally is a diagnostic, not part of the trajectory)
        for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
        if ((threadIdx.x & 31) == 0) atomicAdd(sum_bdenergy, e);
        }
    }

__global__ void k_saru_draw(unsigned int* __restrict__ u32_out, float* __restrict__ f_out,
                            const unsigned int* __restrict__ seeds3, unsigned int ntriples, unsigned int ndraws)
    {
    unsigned int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntriples) return;
    if (u32_out)
        {
        Saru s(seeds3[3 * t], seeds3[3 * t + 1], seeds3[3 * t + 2]);
        for (unsigned int d = 0; d < ndraws; d++) u32_out[(size_t)t * ndraws + d] = s.u32();
        }
    if (f_out)
        {
        Saru s(seeds3[3 * t], seeds3[3 * t + 1], seeds3[3 * t + 2]);
        for (unsigned int d = 0; d < ndraws; d++) f_out[(size_t)t * ndraws + d] = s.s(-1.0f, 1.0f);
        }
    }
} // namespace b200

// ---- ComputeThermo (hoomd/ComputeThermo.cc:195-380): double-precision sums, fixed two-pass order ------------------------
namespace b200
{
constexpr int TH_NQ = 9; // ke2 (= sum m v^2), pe, virial trace, then the 6 virial components

__global__ void __launch_bounds__(256) k_thermo_partial(double* __restrict__ partial, const float4* __restrict__ vel,
                                                        const float4* __restrict__ net_force,
                                                        const float* __restrict__ net_virial,
                                                        unsigned int virial_pitch,
                                                        const unsigned int* __restrict__ group,
                                                        unsigned int group_size)
    {
    double acc[TH_NQ];
#pragma unroll
    for (int q = 0; q < TH_NQ; q++) acc[q] = 0.0;
    // contiguous chunk per block, fixed stride per thread: the summation order depends only on (grid, block)
    for (unsigned int g = blockIdx.x * blockDim.x + threadIdx.x; g < group_size; g += gridDim.x * blockDim.x)
        {
        const unsigned int j = group ? group[g] : g;
        const float4 v = vel[j];
        acc[0] += (double)v.w * ((double)v.x * (double)v.x + (double)v.y * (double)v.y + (double)v.z * (double)v.z);
        acc[1] += (double)net_force[j].w;
        if (net_virial)
            {
#pragma unroll
            for (int k = 0; k < 6; k++) acc[3 + k] += (double)net_virial[k * virial_pitch + j];
            }
        }
    acc[2] = acc[3] + acc[6] + acc[8];
    __shared__ double s[8][TH_NQ];
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int q = 0; q < TH_NQ; q++)
        {
        double x = acc[q];
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0) s[warp][q] = x;
        }
    __syncthreads();
    if (threadIdx.x < TH_NQ)
        {
        double x = 0.0;
        for (int w = 0; w < 8; w++) x += s[w][threadIdx.x];
        partial[(size_t)blockIdx.x * TH_NQ + threadIdx.x] = x;
        }
    }

__global__ void k_thermo_final(double* __restrict__ out, const double* __restrict__ partial, unsigned int nblocks)
    {
    if (threadIdx.x < TH_NQ)
        {
        double x = 0.0;
        for (unsigned int b = 0; b < nblocks; b++) x += partial[(size_t)b * TH_NQ + threadIdx.x];
        if (threadIdx.x == 0) x *= 0.5; // KE = 1/2 sum m v^2
        out[threadIdx.x] = x;
        }
    }
} // namespace b200

// ---- TwoStepNVTMTK (hoomd/md/TwoStepNVTMTKGPU.cu:30-212): exp_fac comes from device memory, so the whole update
// (step one -> kinetic energy -> advanceThermostat -> ... -> step two) is enqueued without a host round trip ------------------
namespace b200
{
__global__ void __launch_bounds__(256) k_nvt_step_one(float4* __restrict__ pos, float4* __restrict__ vel,
                                                      const float3s* __restrict__ accel, int3* __restrict__ image,
                                                      const unsigned int* __restrict__ group, unsigned int group_size,
                                                      Box box, const NvtState* __restrict__ state, float dt)
    {
    const unsigned int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= group_size) return;
    const unsigned int j = group ? group[g] : g;
    float4 p = pos[j], v = vel[j];
    int3 img = image[j];
    nvt_step_one_body(p, v, accel[j], img, box, dt, state->exp_fac);
    pos[j] = p;
    vel[j] = v;
    image[j] = img;
    }

__global__ void __launch_bounds__(256) k_nvt_step_two(float4* __restrict__ vel, float3s* __restrict__ accel,
                                                      const unsigned int* __restrict__ group, unsigned int group_size,
                                                      const float4* __restrict__ net_force,
                                                      const NvtState* __restrict__ state, float dt)
    {
    const unsigned int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= group_size) return;
    const unsigned int j = group ? group[g] : g;
    float4 v = vel[j];
    float3s a;
    nvt_step_two_body(v, a, net_force[j], dt, state->exp_fac);
    vel[j] = v;
    accel[j] = a;
    }

// sum(m v^2) over the group: per-CTA partial sums in a fixed order (pass 1 of the thermostat's ComputeThermo)
__global__ void __launch_bounds__(256) k_nvt_ke_partial(double* __restrict__ partial, const float4* __restrict__ vel,
                                                        const unsigned int* __restrict__ group, unsigned int group_size)
    {
    double acc = 0.0;
    for (unsigned int g = blockIdx.x * blockDim.x + threadIdx.x; g < group_size; g += gridDim.x * blockDim.x)
        acc += nvt_mv2(vel[group ? group[g] : g]);
    __shared__ double s[8];
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0)
        {
        double x = 0.0;
        for (int w = 0; w < 8; w++) x += s[w];
        partial[blockIdx.x] = x;
        }
    }

// pass 2 + TwoStepNVTMTK::advanceThermostat, one warp
__global__ void k_nvt_advance(NvtState* __restrict__ state, const double* __restrict__ partial, unsigned int nblocks,
                              unsigned int ndof, float T, float tau, float dt, unsigned int gen)
    {
    double x = 0.0;
    for (unsigned int b = threadIdx.x; b < nblocks; b += 32) x += partial[b];
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (threadIdx.x == 0)
        {
        nvt_advance_body(state, x, ndof, T, tau, dt);
        if (gen != 0xffffffffu) state->gen = gen; // engine: tag for the hand-off to the step kernel
        }
    }

// Domain decomposition: this rank's sum(m v^2) in fixed point into slot 0 of the accumulators of `parity` (the other slots of
// that parity zeroed); the step barrier adds up the ranks, md_step_nvt_begin advances the thermostat (fused_step.cu)
__global__ void k_nvt_seed(unsigned long long* __restrict__ acc, const double* __restrict__ partial, unsigned int nblocks,
                           double scale, unsigned int parity)
    {
    double x = 0.0;
    for (unsigned int b = threadIdx.x; b < nblocks; b += 32) x += partial[b];
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    unsigned long long* a = acc + (size_t)parity * NVT_SLOTS * NVT_SLOT_STRIDE;
    if (threadIdx.x < NVT_SLOTS)
        a[threadIdx.x * NVT_SLOT_STRIDE] = threadIdx.x == 0 ? (unsigned long long)__double2ll_rn(x * scale) : 0ull;
    }

cudaError_t launch_nvt_seed(unsigned long long* d_acc, double* d_partial, unsigned int n_partial_blocks, const float* d_vel,
                            const unsigned int* d_group, unsigned int group_size, double scale, unsigned int parity,
                            cudaStream_t stream)
    {
    if (n_partial_blocks == 0 || !d_partial || !d_acc) return cudaErrorInvalidValue;
    k_nvt_ke_partial<<<n_partial_blocks, 256, 0, stream>>>(d_partial, (const float4*)d_vel, d_group, group_size);
    k_nvt_seed<<<1, 32, 0, stream>>>(d_acc, d_partial, n_partial_blocks, scale, parity);
    return cudaGetLastError();
    }

cudaError_t launch_nvt_advance(float* d_nvt_state, double* d_partial, unsigned int n_partial_blocks, const float* d_vel,
                               const unsigned int* d_group, unsigned int group_size, unsigned int ndof, float T, float tau,
                               float dt, unsigned int gen, cudaStream_t stream)
    {
    if (n_partial_blocks == 0 || !d_partial || !d_nvt_state) return cudaErrorInvalidValue;
    k_nvt_ke_partial<<<n_partial_blocks, 256, 0, stream>>>(d_partial, (const float4*)d_vel, d_group, group_size);
    k_nvt_advance<<<1, 32, 0, stream>>>((NvtState*)d_nvt_state, d_partial, n_partial_blocks, ndof, T, tau, dt, gen);
    return cudaGetLastError();
    }
} // namespace b200

// ---- BoxResizeUpdater::update, particle part (hoomd/BoxResizeUpdater.cc:98-131) --------------------------------------------
namespace b200
{
__global__ void __launch_bounds__(256) k_box_resize(float4* __restrict__ pos, int3* __restrict__ image, unsigned int N,
                                                    Box old_box, Box new_box, int scale_particles)
    {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    float4 p = pos[i];
    if (scale_particles)
        {
        // fractional coordinates in the old box, Cartesian coordinates in the new one (BoxDim.h:238-245,257-263)
        const float3 f = make_fraction_exact(make_float3(p.x, p.y, p.z), old_box, make_float3(0.0f, 0.0f, 0.0f));
        p.x = __fadd_rn(new_box.lo.x, __fmul_rn(f.x, new_box.L.x));
        p.y = __fadd_rn(new_box.lo.y, __fmul_rn(f.y, new_box.L.y));
        p.z = __fadd_rn(new_box.lo.z, __fmul_rn(f.z, new_box.L.z));
        }
    else
        {
        int3 img = image[i];
        box_wrap(p.x, p.y, p.z, img, new_box);
        image[i] = img;
        }
    pos[i] = p;
    }
} // namespace b200

using namespace b200;

extern "C" {

int b200_integrator_sum_net_force(float* d_net_force, float* d_net_virial, unsigned int net_virial_pitch,
                                  const b200_force_list* fl, unsigned int n, int clear, int compute_virial,
                                  b200_stream_t stream)
    {
    if (n == 0) return 0;
    if (compute_virial)
        k_sum_net_force<true><<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>((float4*)d_net_force, d_net_virial,
                                                                                 net_virial_pitch, *fl, n, clear);
    else
        k_sum_net_force<false><<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>((float4*)d_net_force, d_net_virial,
                                                                                  net_virial_pitch, *fl, n, clear);
    return (int)cudaGetLastError();
    }

int b200_nve_step_one(float* d_pos, float* d_vel, const float* d_accel, int* d_image,
                      const unsigned int* d_group_members, unsigned int group_size, const b200_box* box, float deltaT,
                      int limit, float limit_val, int zero_force, b200_stream_t stream)
    {
    if (!box_is_orthorhombic(*box)) return (int)cudaErrorNotSupported;
    if (group_size == 0) return 0;
    k_nve_step_one<<<(group_size + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        (float4*)d_pos, (float4*)d_vel, (const float3s*)d_accel, (int3*)d_image, d_group_members, group_size,
        make_box(*box), deltaT, limit != 0, limit_val, zero_force != 0);
    return (int)cudaGetLastError();
    }

int b200_nve_step_two(float* d_vel, float* d_accel, const unsigned int* d_group_members, unsigned int group_size,
                      const float* d_net_force, float deltaT, int limit, float limit_val, int zero_force,
                      b200_stream_t stream)
    {
    if (group_size == 0) return 0;
    k_nve_step_two<<<(group_size + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        (float4*)d_vel, (float3s*)d_accel, d_group_members, group_size, (const float4*)d_net_force, deltaT, limit != 0,
        limit_val, zero_force != 0);
    return (int)cudaGetLastError();
    }

unsigned int b200_langevin_hash_seed(unsigned int s)
    {
    s = s * 0x12345677u + 0x12345u;
    s ^= (s >> 16);
    s *= 0x45679u;
    return s;
    }

int b200_langevin_step_two(const float* d_pos, float* d_vel, float* d_accel, const float* d_diameter,
                           const unsigned int* d_tag, const unsigned int* d_group_members, unsigned int group_size,
                           const float* d_net_force, const b200_langevin_args* a, float deltaT, unsigned int D,
                           b200_stream_t stream_)
    {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (group_size == 0) return 0;
    if (a->use_lambda && !d_diameter) return (int)cudaErrorInvalidValue;
    if (a->tally)
        {
        if (!a->d_sum_bdenergy) return (int)cudaErrorInvalidValue;
        cudaMemsetAsync(a->d_sum_bdenergy, 0, sizeof(float), stream);
        }
    k_langevin_step_two<<<(group_size + 255) / 256, 256, 0, stream>>>(
        (const float4*)d_pos, (float4*)d_vel, (float3s*)d_accel, d_diameter, d_tag, d_group_members, group_size,
        (const float4*)d_net_force, a->d_gamma, a->use_lambda != 0, a->lambda, a->T, a->timestep, a->seed,
        a->d_sum_bdenergy, a->noiseless_t != 0, a->tally != 0, deltaT, D);
    return (int)cudaGetLastError();
    }

int b200_nvt_mtk_step_one(float* d_pos, float* d_vel, const float* d_accel, int* d_image,
                          const unsigned int* d_group_members, unsigned int group_size, const b200_box* box,
                          const float* d_nvt_state, float deltaT, b200_stream_t stream)
    {
    if (group_size == 0) return 0;
    if (!box_is_orthorhombic(*box)) return (int)cudaErrorNotSupported;
    k_nvt_step_one<<<(group_size + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        (float4*)d_pos, (float4*)d_vel, (const float3s*)d_accel, (int3*)d_image, d_group_members, group_size,
        make_box(*box), (const NvtState*)d_nvt_state, deltaT);
    return (int)cudaGetLastError();
    }

int b200_nvt_mtk_step_two(float* d_vel, float* d_accel, const unsigned int* d_group_members, unsigned int group_size,
                          const float* d_net_force, const float* d_nvt_state, float deltaT, b200_stream_t stream)
    {
    if (group_size == 0) return 0;
    k_nvt_step_two<<<(group_size + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        (float4*)d_vel, (float3s*)d_accel, d_group_members, group_size, (const float4*)d_net_force,
        (const NvtState*)d_nvt_state, deltaT);
    return (int)cudaGetLastError();
    }

int b200_nvt_mtk_advance(float* d_nvt_state, double* d_partial, unsigned int n_partial_blocks, const float* d_vel,
                         const unsigned int* d_group_members, unsigned int group_size, unsigned int ndof, float T,
                         float tau, float deltaT, b200_stream_t stream)
    {
    return (int)launch_nvt_advance(d_nvt_state, d_partial, n_partial_blocks, d_vel, d_group_members, group_size, ndof, T,
                                   tau, deltaT, 0xffffffffu, (cudaStream_t)stream);
    }

int b200_box_resize_particles(float* d_pos, int* d_image, unsigned int N, const b200_box* old_box, const b200_box* new_box,
                              int scale_particles, b200_stream_t stream)
    {
    if (N == 0) return 0;
    if (!box_is_orthorhombic(*old_box) || !box_is_orthorhombic(*new_box)) return (int)cudaErrorNotSupported;
    k_box_resize<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>((float4*)d_pos, (int3*)d_image, N, make_box(*old_box),
                                                                     make_box(*new_box), scale_particles);
    return (int)cudaGetLastError();
    }

int b200_saru_draw(unsigned int* d_u32_out, float* d_f_out, const unsigned int* d_seeds3, unsigned int ntriples,
                   unsigned int ndraws, b200_stream_t stream)
    {
    if (ntriples == 0) return 0;
    k_saru_draw<<<(ntriples + 127) / 128, 128, 0, (cudaStream_t)stream>>>(d_u32_out, d_f_out, d_seeds3, ntriples,
                                                                         ndraws);
    return (int)cudaGetLastError();
    }

int b200_compute_thermo(double* d_out, double* d_partial, unsigned int n_partial_blocks, const float* d_vel,
                        const float* d_net_force, const float* d_net_virial, unsigned int virial_pitch,
                        const unsigned int* d_group_members, unsigned int group_size, b200_stream_t stream_)
    {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (n_partial_blocks == 0) return (int)cudaErrorInvalidValue;
    k_thermo_partial<<<n_partial_blocks, 256, 0, stream>>>(d_partial, (const float4*)d_vel, (const float4*)d_net_force,
                                                          d_net_virial, virial_pitch, d_group_members, group_size);
    k_thermo_final<<<1, 32, 0, stream>>>(d_out, d_partial, n_partial_blocks);
    return (int)cudaGetLastError();
    }

} // extern "C"

// ---- ComputeThermo without any allocation (no counterpart in the reference; see include/hoomd_b200.h) ---------------------
// The per-CTA partial sums of k_thermo_partial live in a file-scope __device__ array instead of a caller buffer.  In host
// code the array's NAME is only a stand-in for the symbol: the device address is looked up once per device with
// cudaGetSymbolAddress (that device current) and cached; the kernels get the looked-up pointer.
namespace b200
{
static __device__ __align__(256) double g_thermo_static_scratch[B200_THERMO_STATIC_MAX_BLOCKS * TH_NQ];
constexpr int THERMO_STATIC_MAX_DEVICES = 64;
static double* g_thermo_static_ptr[THERMO_STATIC_MAX_DEVICES] = {nullptr};
static std::mutex g_thermo_static_mutex;

// device address of the scratch array on the CURRENT device (set-up: first call per device)
static cudaError_t thermo_static_scratch(double** out)
    {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= THERMO_STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_thermo_static_mutex);
    if (!g_thermo_static_ptr[dev])
        {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_thermo_static_scratch);
        if (e != cudaSuccess) return e;
        g_thermo_static_ptr[dev] = (double*)p;
        }
    *out = g_thermo_static_ptr[dev];
    return cudaSuccess;
    }
} // namespace b200

extern "C" int b200_compute_thermo_static(double* d_out, unsigned int n_partial_blocks, const float* d_vel,
                                          const float* d_net_force, const float* d_net_virial, unsigned int virial_pitch,
                                          const unsigned int* d_group_members, unsigned int group_size,
                                          b200_stream_t stream_)
    {
    using namespace b200;
    cudaStream_t stream = (cudaStream_t)stream_;
    if (n_partial_blocks == 0 || n_partial_blocks > B200_THERMO_STATIC_MAX_BLOCKS)
        {
        fprintf(stderr, "b200_compute_thermo_static: n_partial_blocks = %u is outside [1, %u] (the built-in scratch array "
                        "holds %u partial sums); use b200_compute_thermo with a caller buffer for more\n",
                n_partial_blocks, (unsigned int)B200_THERMO_STATIC_MAX_BLOCKS, (unsigned int)B200_THERMO_STATIC_MAX_BLOCKS);
        return (int)cudaErrorInvalidValue;
        }
    double* scratch = nullptr;
    const cudaError_t e = thermo_static_scratch(&scratch);
    if (e != cudaSuccess)
        {
        fprintf(stderr, "b200_compute_thermo_static: cudaGetSymbolAddress of the scratch array failed: %s\n",
                cudaGetErrorString(e));
        return (int)e;
        }
    // every partial row [0, n_partial_blocks) is overwritten by pass 1 before pass 2 reads it: nothing has to start at zero,
    // rows beyond n_partial_blocks (left by an earlier, larger call) are never read
    k_thermo_partial<<<n_partial_blocks, 256, 0, stream>>>(scratch, (const float4*)d_vel, (const float4*)d_net_force,
                                                          d_net_virial, virial_pitch, d_group_members, group_size);
    k_thermo_final<<<1, 32, 0, stream>>>(d_out, scratch, n_partial_blocks);
    return (int)cudaGetLastError();
    }

// ---- TwoStepNPTMTK (hoomd/md/TwoStepNPTMTK.cc:177-330,471-533, hoomd/md/TwoStepNPTMTKGPU.cu): the particle loops of the
// Martyna-Tobias-Klein barostat/thermostat.  The propagator matrices (upper triangular, {xx, xy, xz, yy, yz, zz}) and the
// thermostat factor are computed by the caller on the host, exactly as the reference does (updatePropagator); the kernels
// restate the per-particle arithmetic operation by operation (no FMA contraction) so that trajectories follow the
// reference's to the last bit of every half step.
namespace b200
{
struct NptMat { float v[6], r[6], ri[6]; };

// a.x*m0 + a.y*m1 + a.z*m2 evaluated left to right like the C++ expression m0*x + m1*y + m2*z
__device__ __forceinline__ float npt_row3(float m0, float x, float m1, float y, float m2, float z)
    {
    return __fadd_rn(__fadd_rn(__fmul_rn(m0, x), __fmul_rn(m1, y)), __fmul_rn(m2, z));
    }
__device__ __forceinline__ float npt_row2(float m0, float x, float m1, float y)
    {
    return __fadd_rn(__fmul_rn(m0, x), __fmul_rn(m1, y));
    }

__global__ void __launch_bounds__(256) k_npt_step_one(float4* __restrict__ pos, float4* __restrict__ vel,
                                                      const float3s* __restrict__ accel,
                                                      const unsigned int* __restrict__ group, unsigned int group_size,
                                                      const NptMat m, float exp_thermo_fac, float dt, int rescale_all)
    {
    const unsigned int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= group_size) return;
    const unsigned int j = group ? group[g] : g;
    float4 p = pos[j], v = vel[j];
    const float3s a = accel[j];
    // v += deltaT/2 * accel
    const float hdt = __fdiv_rn(dt, 2.0f);
    v.x = __fadd_rn(v.x, __fmul_rn(hdt, a.x));
    v.y = __fadd_rn(v.y, __fmul_rn(hdt, a.y));
    v.z = __fadd_rn(v.z, __fmul_rn(hdt, a.z));
    // barostat: multiply with the matrix exponential (each row reads the not yet overwritten components)
    v.x = npt_row3(m.v[0], v.x, m.v[1], v.y, m.v[2], v.z);
    v.y = npt_row2(m.v[3], v.y, m.v[4], v.z);
    v.z = __fmul_rn(m.v[5], v.z);
    // thermostat
    v.x = __fmul_rn(v.x, exp_thermo_fac);
    v.y = __fmul_rn(v.y, exp_thermo_fac);
    v.z = __fmul_rn(v.z, exp_thermo_fac);
    if (!rescale_all)
        {
        p.x = npt_row3(m.r[0], p.x, m.r[1], p.y, m.r[2], p.z);
        p.y = npt_row2(m.r[3], p.y, m.r[4], p.z);
        p.z = __fmul_rn(m.r[5], p.z);
        }
    p.x = __fadd_rn(p.x, npt_row3(m.ri[0], v.x, m.ri[1], v.y, m.ri[2], v.z));
    p.y = __fadd_rn(p.y, npt_row2(m.ri[3], v.y, m.ri[4], v.z));
    p.z = __fadd_rn(p.z, __fmul_rn(m.ri[5], v.z));
    pos[j] = p;
    vel[j] = v;
    }

__global__ void __launch_bounds__(256) k_npt_rescale(float4* __restrict__ pos, unsigned int N, float rxx, float rxy,
                                                     float rxz, float ryy, float ryz, float rzz)
    {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    float4 p = pos[i];
    p.x = npt_row3(rxx, p.x, rxy, p.y, rxz, p.z);
    p.y = npt_row2(ryy, p.y, ryz, p.z);
    p.z = __fmul_rn(rzz, p.z);
    pos[i] = p;
    }

__global__ void __launch_bounds__(256) k_npt_wrap(float4* __restrict__ pos, int3* __restrict__ image, unsigned int N,
                                                  Box box)
    {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    float4 p = pos[i];
    int3 img = image[i];
    box_wrap(p.x, p.y, p.z, img, box);
    pos[i] = p;
    image[i] = img;
    }

__global__ void __launch_bounds__(256) k_npt_step_two(float4* __restrict__ vel, float3s* __restrict__ accel,
                                                      const unsigned int* __restrict__ group, unsigned int group_size,
                                                      const float4* __restrict__ net_force, const NptMat m, float dt,
                                                      float exp_thermo_fac)
    {
    const unsigned int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= group_size) return;
    const unsigned int j = group ? group[g] : g;
    float4 v = vel[j];
    const float4 f = net_force[j];
    const float minv = __fdiv_rn(1.0f, v.w);
    float3s a;
    a.x = __fmul_rn(f.x, minv);
    a.y = __fmul_rn(f.y, minv);
    a.z = __fmul_rn(f.z, minv);
    v.x = __fmul_rn(v.x, exp_thermo_fac);
    v.y = __fmul_rn(v.y, exp_thermo_fac);
    v.z = __fmul_rn(v.z, exp_thermo_fac);
    v.x = npt_row3(m.v[0], v.x, m.v[1], v.y, m.v[2], v.z);
    v.y = npt_row2(m.v[3], v.y, m.v[4], v.z);
    v.z = __fmul_rn(m.v[5], v.z);
    const float hdt = __fdiv_rn(dt, 2.0f);
    v.x = __fadd_rn(v.x, __fmul_rn(hdt, a.x));
    v.y = __fadd_rn(v.y, __fmul_rn(hdt, a.y));
    v.z = __fadd_rn(v.z, __fmul_rn(hdt, a.z));
    vel[j] = v;
    accel[j] = a;
    }

// ComputeThermo with pdata_flag::pressure_tensor (hoomd/ComputeThermo.cc:224-240,300-318): the six components of
// sum(m v v) and of the virial, and the potential energy, in double
constexpr int TT_NQ = 13;
__global__ void __launch_bounds__(256) k_thermo_tensor_partial(double* __restrict__ partial, const float4* __restrict__ vel,
                                                               const float4* __restrict__ net_force,
                                                               const float* __restrict__ net_virial,
                                                               unsigned int virial_pitch,
                                                               const unsigned int* __restrict__ group,
                                                               unsigned int group_size)
    {
    double acc[TT_NQ];
#pragma unroll
    for (int q = 0; q < TT_NQ; q++) acc[q] = 0.0;
    for (unsigned int g = blockIdx.x * blockDim.x + threadIdx.x; g < group_size; g += gridDim.x * blockDim.x)
        {
        const unsigned int j = group ? group[g] : g;
        const float4 v = vel[j];
        const double mass = (double)v.w, x = (double)v.x, y = (double)v.y, z = (double)v.z;
        acc[0] += mass * (x * x);
        acc[1] += mass * (x * y);
        acc[2] += mass * (x * z);
        acc[3] += mass * (y * y);
        acc[4] += mass * (y * z);
        acc[5] += mass * (z * z);
        if (net_force) acc[6] += (double)net_force[j].w;
        if (net_virial)
            {
#pragma unroll
            for (int k = 0; k < 6; k++) acc[7 + k] += (double)net_virial[k * virial_pitch + j];
            }
        }
    __shared__ double s[8][TT_NQ];
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int q = 0; q < TT_NQ; q++)
        {
        double x = acc[q];
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0) s[warp][q] = x;
        }
    __syncthreads();
    if (threadIdx.x < TT_NQ)
        {
        double x = 0.0;
        for (int w = 0; w < 8; w++) x += s[w][threadIdx.x];
        partial[(size_t)blockIdx.x * TT_NQ + threadIdx.x] = x;
        }
    }

__global__ void k_thermo_tensor_final(double* __restrict__ out, const double* __restrict__ partial, unsigned int nblocks)
    {
    if (threadIdx.x < TT_NQ)
        {
        double x = 0.0;
        for (unsigned int b = 0; b < nblocks; b++) x += partial[(size_t)b * TT_NQ + threadIdx.x];
        out[threadIdx.x] = x;
        }
    }
} // namespace b200

extern "C" {

int b200_npt_mtk_step_one(float* d_pos, float* d_vel, const float* d_accel, const unsigned int* d_group_members,
                          unsigned int group_size, float exp_thermo_fac, const float* mat_exp_v, const float* mat_exp_r,
                          const float* mat_exp_r_int, float deltaT, int rescale_all, b200_stream_t stream)
    {
    if (group_size == 0) return 0;
    NptMat m;
    for (int k = 0; k < 6; k++)
        {
        m.v[k] = mat_exp_v[k];
        m.r[k] = mat_exp_r[k];
        m.ri[k] = mat_exp_r_int[k];
        }
    k_npt_step_one<<<(group_size + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        (float4*)d_pos, (float4*)d_vel, (const float3s*)d_accel, d_group_members, group_size, m, exp_thermo_fac, deltaT,
        rescale_all);
    return (int)cudaGetLastError();
    }

int b200_npt_mtk_wrap(unsigned int N, float* d_pos, int* d_image, const b200_box* box, b200_stream_t stream)
    {
    if (N == 0) return 0;
    if (!box_is_orthorhombic(*box)) return (int)cudaErrorNotSupported;
    k_npt_wrap<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>((float4*)d_pos, (int3*)d_image, N, make_box(*box));
    return (int)cudaGetLastError();
    }

int b200_npt_mtk_step_two(float* d_vel, float* d_accel, const unsigned int* d_group_members, unsigned int group_size,
                          const float* d_net_force, const float* mat_exp_v, float deltaT, float exp_thermo_fac,
                          b200_stream_t stream)
    {
    if (group_size == 0) return 0;
    NptMat m;
    for (int k = 0; k < 6; k++)
        {
        m.v[k] = mat_exp_v[k];
        m.r[k] = m.ri[k] = 0.0f;
        }
    k_npt_step_two<<<(group_size + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        (float4*)d_vel, (float3s*)d_accel, d_group_members, group_size, (const float4*)d_net_force, m, deltaT,
        exp_thermo_fac);
    return (int)cudaGetLastError();
    }

int b200_npt_mtk_rescale(unsigned int N, float* d_pos, float mat_exp_r_xx, float mat_exp_r_xy, float mat_exp_r_xz,
                         float mat_exp_r_yy, float mat_exp_r_yz, float mat_exp_r_zz, b200_stream_t stream)
    {
    if (N == 0) return 0;
    k_npt_rescale<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>((float4*)d_pos, N, mat_exp_r_xx, mat_exp_r_xy,
                                                                    mat_exp_r_xz, mat_exp_r_yy, mat_exp_r_yz,
                                                                    mat_exp_r_zz);
    return (int)cudaGetLastError();
    }

int b200_compute_thermo_tensor(double* d_out, double* d_partial, unsigned int n_partial_blocks, const float* d_vel,
                               const float* d_net_force, const float* d_net_virial, unsigned int virial_pitch,
                               const unsigned int* d_group_members, unsigned int group_size, b200_stream_t stream_)
    {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (n_partial_blocks == 0 || !d_partial || !d_out) return (int)cudaErrorInvalidValue;
    k_thermo_tensor_partial<<<n_partial_blocks, 256, 0, stream>>>(d_partial, (const float4*)d_vel,
                                                                 (const float4*)d_net_force, d_net_virial, virial_pitch,
                                                                 d_group_members, group_size);
    k_thermo_tensor_final<<<1, 32, 0, stream>>>(d_out, d_partial, n_partial_blocks);
    return (int)cudaGetLastError();
    }

} // extern "C"
