// Per-particle bodies of the integration half steps, shared by the driver kernels (integrate.cu) and the fused
// device-driven step kernels (fused_step.cu) so both produce bit-identical results.
#pragma once
#include "common.cuh"

namespace b200
{
struct float3s { float x, y, z; }; // Scalar3 with 4 B alignment, the reference's accel layout

// ---- TwoStepNVE::integrateStepOne ------------------------------------------------------------------------------------
__device__ __forceinline__ void nve_step_one_body(float4& p, float4& v, float3s a, int3& img, const Box& box,
                                                  float dt, bool limit, float limit_val)
    {
    const float hx = __fmul_rn(0.5f, a.x), hy = __fmul_rn(0.5f, a.y), hz = __fmul_rn(0.5f, a.z);
    // dx = v*dt + (1/2)*a*dt*dt, evaluated left to right
    float dx = __fadd_rn(__fmul_rn(v.x, dt), __fmul_rn(__fmul_rn(hx, dt), dt));
    float dy = __fadd_rn(__fmul_rn(v.y, dt), __fmul_rn(__fmul_rn(hy, dt), dt));
    float dz = __fadd_rn(__fmul_rn(v.z, dt), __fmul_rn(__fmul_rn(hz, dt), dt));
    if (limit)
        {
        const float len = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
        if (len > limit_val)
            {
            dx = __fmul_rn(__fdiv_rn(dx, len), limit_val);
            dy = __fmul_rn(__fdiv_rn(dy, len), limit_val);
            dz = __fmul_rn(__fdiv_rn(dz, len), limit_val);
            }
        }
    p.x = __fadd_rn(p.x, dx);
    p.y = __fadd_rn(p.y, dy);
    p.z = __fadd_rn(p.z, dz);
    v.x = __fadd_rn(v.x, __fmul_rn(hx, dt));
    v.y = __fadd_rn(v.y, __fmul_rn(hy, dt));
    v.z = __fadd_rn(v.z, __fmul_rn(hz, dt));
    box_wrap(p.x, p.y, p.z, img, box);
    }

// ---- TwoStepNVE::integrateStepTwo ------------------------------------------------------------------------------------
__device__ __forceinline__ void nve_step_two_body(float4& v, float3s& a, float4 f, float dt, bool limit,
                                                  float limit_val, bool zero_force)
    {
    if (zero_force)
        a.x = a.y = a.z = 0.0f;
    else
        {
        const float minv = __frcp_rn(v.w); // == Scalar(1.0)/mass, correctly rounded
        a.x = __fmul_rn(f.x, minv);
        a.y = __fmul_rn(f.y, minv);
        a.z = __fmul_rn(f.z, minv);
        }
    v.x = __fadd_rn(v.x, __fmul_rn(__fmul_rn(0.5f, a.x), dt));
    v.y = __fadd_rn(v.y, __fmul_rn(__fmul_rn(0.5f, a.y), dt));
    v.z = __fadd_rn(v.z, __fmul_rn(__fmul_rn(0.5f, a.z), dt));
    if (limit)
        {
        const float s = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(v.x, v.x), __fmul_rn(v.y, v.y)), __fmul_rn(v.z, v.z)));
        if (__fmul_rn(s, dt) > limit_val)
            {
            v.x = __fdiv_rn(__fmul_rn(__fdiv_rn(v.x, s), limit_val), dt);
            v.y = __fdiv_rn(__fmul_rn(__fdiv_rn(v.y, s), limit_val), dt);
            v.z = __fdiv_rn(__fmul_rn(__fdiv_rn(v.z, s), limit_val), dt);
            }
        }
    }

// ---- TwoStepLangevin::integrateStepTwo -------------------------------------------------------------------------------
__device__ __forceinline__ float langevin_step_two_body(float4& v, float3s& a, float4 f, unsigned int ptag,
                                                        unsigned int timestep, unsigned int seed, float gamma,
                                                        float T, float dt, bool noiseless_t, unsigned int D)
    {
    Saru saru(ptag, timestep, seed);
    const float rx = saru.s(-1.0f, 1.0f);
    const float ry = saru.s(-1.0f, 1.0f);
    const float rz = saru.s(-1.0f, 1.0f);
    // coeff = sqrt(6*gamma*T/dt), left to right
    float coeff = __fsqrt_rn(__fdiv_rn(__fmul_rn(__fmul_rn(6.0f, gamma), T), dt));
    if (noiseless_t) coeff = 0.0f;
    const float bd_fx = __fsub_rn(__fmul_rn(rx, coeff), __fmul_rn(gamma, v.x));
    const float bd_fy = __fsub_rn(__fmul_rn(ry, coeff), __fmul_rn(gamma, v.y));
    float bd_fz = __fsub_rn(__fmul_rn(rz, coeff), __fmul_rn(gamma, v.z));
    if (D < 3) bd_fz = 0.0f;
    const float minv = __frcp_rn(v.w); // == Scalar(1.0)/mass, correctly rounded
    a.x = __fmul_rn(__fadd_rn(f.x, bd_fx), minv);
    a.y = __fmul_rn(__fadd_rn(f.y, bd_fy), minv);
    a.z = __fmul_rn(__fadd_rn(f.z, bd_fz), minv);
    v.x = __fadd_rn(v.x, __fmul_rn(__fmul_rn(0.5f, a.x), dt));
    v.y = __fadd_rn(v.y, __fmul_rn(__fmul_rn(0.5f, a.y), dt));
    v.z = __fadd_rn(v.z, __fmul_rn(__fmul_rn(0.5f, a.z), dt));
    // reservoir energy transfer (hoomd/md/TwoStepLangevin.cc:306)
    return __fadd_rn(__fadd_rn(__fmul_rn(bd_fx, v.x), __fmul_rn(bd_fy, v.y)), __fmul_rn(bd_fz, v.z));
    }

} // namespace b200
