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

// ---- TwoStepNVTMTK, point particles (hoomd/md/TwoStepNVTMTK.cc:107-160,297-340,387-404) ------------------------------------
// Thermostat state on the device: IntegratorVariables xi, eta and the loop-invariant m_exp_thermo_fac.  Layout of the
// `float d_nvt_state[4]` of the C ABI.
struct NvtState { float xi, eta, exp_fac; unsigned int gen; }; // gen: engine's hand-off tag (ts of the last advance)

__device__ __forceinline__ void nvt_step_one_body(float4& p, float4& v, float3s a, int3& img, const Box& box, float dt,
                                                  float exp_fac)
    {
    // v = (v + (1/2)*accel*dt) * exp_fac ; pos += dt*v ; wrap
    v.x = __fmul_rn(__fadd_rn(v.x, __fmul_rn(__fmul_rn(0.5f, a.x), dt)), exp_fac);
    v.y = __fmul_rn(__fadd_rn(v.y, __fmul_rn(__fmul_rn(0.5f, a.y), dt)), exp_fac);
    v.z = __fmul_rn(__fadd_rn(v.z, __fmul_rn(__fmul_rn(0.5f, a.z), dt)), exp_fac);
    p.x = __fadd_rn(p.x, __fmul_rn(dt, v.x));
    p.y = __fadd_rn(p.y, __fmul_rn(dt, v.y));
    p.z = __fadd_rn(p.z, __fmul_rn(dt, v.z));
    box_wrap(p.x, p.y, p.z, img, box);
    }

__device__ __forceinline__ void nvt_step_two_body(float4& v, float3s& a, float4 f, float dt, float exp_fac)
    {
    const float minv = __frcp_rn(v.w);
    a.x = __fmul_rn(f.x, minv);
    a.y = __fmul_rn(f.y, minv);
    a.z = __fmul_rn(f.z, minv);
    // v = v*exp_fac + (1/2)*dt*accel, the scalar product first
    const float hdt = __fmul_rn(0.5f, dt);
    v.x = __fadd_rn(__fmul_rn(v.x, exp_fac), __fmul_rn(hdt, a.x));
    v.y = __fadd_rn(__fmul_rn(v.y, exp_fac), __fmul_rn(hdt, a.y));
    v.z = __fadd_rn(__fmul_rn(v.z, exp_fac), __fmul_rn(hdt, a.z));
    }

// m v^2 of one particle in double, the summand of ComputeThermo's kinetic energy (hoomd/ComputeThermo.cc:241-252)
__device__ __forceinline__ double nvt_mv2(const float4 v)
    {
    return (double)v.w * ((double)v.x * (double)v.x + (double)v.y * (double)v.y + (double)v.z * (double)v.z);
    }

// TwoStepNVTMTK::advanceThermostat from sum(m v^2) of the half-step velocities.  m_T->getValue() is a double in the
// reference, so the bracket and the two sums are evaluated in double and stored as Scalar; no FMA contraction.
__device__ __forceinline__ void nvt_advance_body(NvtState* st, double mv2_sum, unsigned int ndof, float T, float tau,
                                                 float dt)
    {
    const float ke = (float)__dmul_rn(mv2_sum, 0.5);
    const float curr_T = __fmul_rn(__fdiv_rn(2.0f, (float)ndof), ke); // ComputeThermo::getTranslationalTemperature
    const double dev = __dsub_rn(__ddiv_rn((double)curr_T, (double)T), 1.0);
    const float c = __fdiv_rn(__fdiv_rn(__fmul_rn(0.5f, dt), tau), tau);
    const double step = __dmul_rn((double)c, dev);
    const float xi_prime = (float)__dadd_rn((double)st->xi, step);
    const float xi = (float)__dadd_rn((double)xi_prime, step);
    st->xi = xi;
    st->eta = __fadd_rn(st->eta, __fmul_rn(xi_prime, dt));
    // exp(-(1/2)*xi*dt) as the host libm's expf gives it: round the double result (correctly rounded but for ties)
    st->exp_fac = (float)exp((double)__fmul_rn(__fmul_rn(-0.5f, xi), dt));
    }

} // namespace b200
