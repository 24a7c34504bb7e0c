// Bonded pair forces and the neighbor-list exclusion filter (Kremer-Grest = WCA pair + FENE bonds).
//   k_bond_forces<Evaluator, TAGKEYED> : PotentialBond<evaluator>::computeForces (hoomd/md/PotentialBond.h:155-300), one lane per
//                                        particle walks its bond rows like gpu_compute_bond_forces_kernel
//                                        (hoomd/md/PotentialBondGPU.cuh:112-255)
//   k_nlist_filter<TAGKEYED>           : NeighborList::filterNlist (hoomd/md/NeighborList.cc:1275-1328), one WARP per row:
//                                        coalesced chunk loads, ballot + popc compaction in place
//   k_add_force                        : net_force += force (Integrator::computeNetForce's loop over the ForceComputes,
//                                        hoomd/Integrator.cc:308-420) for the stand-alone force evaluation
// TAGKEYED tables are indexed by the particle TAG and hold partner TAGS, resolved through rtag at use: the tables never
// have to be rebuilt when the SFC sorter permutes the particles (the reference rebuilds its index-keyed GPU tables,
// hoomd/BondedGroupData.cc rebuildGPUTable, and its m_ex_list_idx, NeighborList.cc:1234-1270, after every sort).
#include "common.cuh"
#include "kernels.h"

namespace b200
{
// EvaluatorBondFENE::evalForceAndEnergy (hoomd/md/EvaluatorBondFENE.h:95-133); needsDiameter() == true
struct EvaluatorBondFENE
    {
    typedef float4 param_type; // (K, r_0, lj1, lj2)
    __device__ __forceinline__ static bool eval(float rsq, float da, float db, const float4& p, float& force_divr,
                                                float& bond_eng)
        {
        const float K = p.x, r_0 = p.y, lj1 = p.z, lj2 = p.w;
        const float r = __fsqrt_rn(rsq);
        // sqrt(rsq) - diameter_a/2 - diameter_b/2 + 1.0, left to right
        const float rtemp = __fadd_rn(__fsub_rn(__fsub_rn(r, __fdiv_rn(da, 2.0f)), __fdiv_rn(db, 2.0f)), 1.0f);
        const float rmdoverr = __fdiv_rn(rtemp, r);
        rsq = __fmul_rn(rtemp, rtemp);
        const float r2inv = __fdiv_rn(1.0f, rsq);
        const float r6inv = __fmul_rn(__fmul_rn(r2inv, r2inv), r2inv);
        float wca_divr = 0.0f, pair_eng = 0.0f;
        const float sigma6inv = __fdiv_rn(lj2, lj1);
        const float epsilon = __fdiv_rn(__fdiv_rn(__fmul_rn(lj2, lj2), 4.0f), lj1);
        if (r6inv > __fdiv_rn(sigma6inv, 2.0f)) // wcalimit (2^(1/6))^6 sigma^6
            {
            wca_divr = __fmul_rn(__fmul_rn(r2inv, r6inv),
                                 __fsub_rn(__fmul_rn(__fmul_rn(12.0f, lj1), r6inv), __fmul_rn(6.0f, lj2)));
            pair_eng = __fadd_rn(__fmul_rn(r6inv, __fsub_rn(__fmul_rn(lj1, r6inv), lj2)), epsilon);
            }
        force_divr = __fmul_rn(wca_divr, rmdoverr);
        bond_eng = pair_eng;
        if (K != 0.0f)
            {
            const float r0sq = __fmul_rn(r_0, r_0);
            if (rsq >= r0sq) return false;
            const float q = __fdiv_rn(rsq, r0sq);
            force_divr = __fadd_rn(force_divr, __fmul_rn(__fdiv_rn(-K, __fsub_rn(1.0f, q)), rmdoverr));
            bond_eng = __fadd_rn(bond_eng, __fmul_rn(__fmul_rn(__fmul_rn(-0.5f, K), r0sq), logf(__fsub_rn(1.0f, q))));
            }
        return true;
        }
    };

// EvaluatorBondHarmonic::evalForceAndEnergy (hoomd/md/EvaluatorBondHarmonic.h:75-92)
struct EvaluatorBondHarmonic
    {
    typedef float2 param_type; // (K, r_0)
    __device__ __forceinline__ static bool eval(float rsq, float, float, const float2& p, float& force_divr,
                                                float& bond_eng)
        {
        const float K = p.x, r_0 = p.y;
        const float r = __fsqrt_rn(rsq);
        force_divr = __fmul_rn(K, __fsub_rn(__fdiv_rn(r_0, r), 1.0f));
        if (!isfinite(force_divr)) force_divr = 0.0f;
        const float d = __fsub_rn(r_0, r);
        bond_eng = __fmul_rn(__fmul_rn(__fmul_rn(0.5f, K), d), d);
        return true;
        }
    };

constexpr unsigned int BOND_MAX_TYPES_SMEM = 256;

template<class Evaluator, bool TAGKEYED>
__global__ void __launch_bounds__(256) k_bond_forces(float4* __restrict__ force, float* __restrict__ virial,
                                                     unsigned int virial_pitch, unsigned int N,
                                                     const float4* __restrict__ pos, const float* __restrict__ diameter,
                                                     Box box, const uint2* __restrict__ blist, unsigned int pitch,
                                                     const unsigned int* __restrict__ n_bonds, unsigned int n_bond_types,
                                                     const typename Evaluator::param_type* __restrict__ params,
                                                     const unsigned int* __restrict__ tag,
                                                     const unsigned int* __restrict__ rtag, bool compute_virial,
                                                     unsigned int* __restrict__ flags)
    {
    __shared__ typename Evaluator::param_type s_params[BOND_MAX_TYPES_SMEM];
    for (unsigned int k = threadIdx.x; k < n_bond_types; k += blockDim.x) s_params[k] = params[k];
    __syncthreads();
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const unsigned int row = TAGKEYED ? tag[i] : i;
    // (domain decomposition: unused local slots carry the tag 0xffffffff)
    const unsigned int nb = (TAGKEYED && row == 0xffffffffu) ? 0u : n_bonds[row];
    const float4 pi4 = pos[i];
    const float di = diameter ? diameter[i] : 1.0f;
    float4 F = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    float vir[6] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    for (unsigned int k = 0; k < nb; k++)
        {
        const uint2 e = blist[(size_t)k * pitch + row];
        const unsigned int j = TAGKEYED ? rtag[e.x] : e.x;
        if (TAGKEYED && j == 0xffffffffu)
            {
            *flags = 2u; // domain decomposition: the partner is neither local nor a ghost ("bond incomplete")
            return;
            }
        const float4 pj4 = pos[j];
        const float dj = diameter ? diameter[j] : 1.0f;
        float3 dx = make_float3(__fsub_rn(pi4.x, pj4.x), __fsub_rn(pi4.y, pj4.y), __fsub_rn(pi4.z, pj4.z));
        dx = min_image_exact(dx, box);
        const float rsq = dot_exact(dx);
        float force_divr = 0.0f, bond_eng = 0.0f;
        if (!Evaluator::eval(rsq, di, dj, s_params[e.y], force_divr, bond_eng))
            {
            *flags = 1u; // "bond out of bounds" (hoomd/md/PotentialBond.h:288-292); the host raises
            return;
            }
        F.x = __fadd_rn(F.x, __fmul_rn(dx.x, force_divr));
        F.y = __fadd_rn(F.y, __fmul_rn(dx.y, force_divr));
        F.z = __fadd_rn(F.z, __fmul_rn(dx.z, force_divr));
        F.w = __fadd_rn(F.w, __fmul_rn(bond_eng, 0.5f));
        if (compute_virial)
            {
            const float h = __fmul_rn(0.5f, force_divr);
            vir[0] = __fadd_rn(vir[0], __fmul_rn(__fmul_rn(dx.x, dx.x), h));
            vir[1] = __fadd_rn(vir[1], __fmul_rn(__fmul_rn(dx.x, dx.y), h));
            vir[2] = __fadd_rn(vir[2], __fmul_rn(__fmul_rn(dx.x, dx.z), h));
            vir[3] = __fadd_rn(vir[3], __fmul_rn(__fmul_rn(dx.y, dx.y), h));
            vir[4] = __fadd_rn(vir[4], __fmul_rn(__fmul_rn(dx.y, dx.z), h));
            vir[5] = __fadd_rn(vir[5], __fmul_rn(__fmul_rn(dx.z, dx.z), h));
            }
        }
    force[i] = F;
    if (compute_virial)
        {
#pragma unroll
        for (int c = 0; c < 6; c++) virial[(size_t)c * virial_pitch + i] = vir[c];
        }
    }

template<class Evaluator>
static cudaError_t bond_forces(const b200_bond_args* a, const float* d_params, unsigned int* d_flags,
                               const unsigned int* enable_unused, cudaStream_t stream)
    {
    (void)enable_unused;
    if (!a || !d_params || !d_flags) return cudaErrorInvalidValue;
    if (a->N == 0) return cudaSuccess;
    if (!box_is_orthorhombic(a->box)) return cudaErrorNotSupported;
    if (a->n_bond_types == 0 || a->n_bond_types > BOND_MAX_TYPES_SMEM) return cudaErrorInvalidValue;
    if ((a->d_tag == nullptr) != (a->d_rtag == nullptr)) return cudaErrorInvalidValue;
    if (a->compute_virial && !a->d_virial) return cudaErrorInvalidValue;
    const unsigned int grid = (a->N + 255) / 256;
    typedef typename Evaluator::param_type P;
    if (a->d_tag)
        k_bond_forces<Evaluator, true><<<grid, 256, 0, stream>>>(
            (float4*)a->d_force, a->d_virial, a->virial_pitch, a->N, (const float4*)a->d_pos, a->d_diameter,
            make_box(a->box), (const uint2*)a->d_bond_list, a->pitch, a->d_n_bonds, a->n_bond_types, (const P*)d_params,
            a->d_tag, a->d_rtag, a->compute_virial != 0, d_flags);
    else
        k_bond_forces<Evaluator, false><<<grid, 256, 0, stream>>>(
            (float4*)a->d_force, a->d_virial, a->virial_pitch, a->N, (const float4*)a->d_pos, a->d_diameter,
            make_box(a->box), (const uint2*)a->d_bond_list, a->pitch, a->d_n_bonds, a->n_bond_types, (const P*)d_params,
            nullptr, nullptr, a->compute_virial != 0, d_flags);
    return cudaGetLastError();
    }

// ---- NeighborList::filterNlist ------------------------------------------------------------------------------------------
constexpr int FILTER_MAX_EX = 8; // exclusions held in registers per pass (the reference's FILTER_BATCH_SIZE is 4)

// G lanes per row (G = 8: four rows per warp for short rows such as the ~12-entry rows of a bead-spring melt; G = 32: one
// row per warp).  Every lane of the warp takes part in every ballot; a row's lanes look at their own G-bit segment.
template<bool TAGKEYED, int G>
__global__ void __launch_bounds__(256) k_nlist_filter(unsigned int* __restrict__ n_neigh, unsigned int* __restrict__ nlist,
                                                      const unsigned int* __restrict__ head_list,
                                                      const unsigned int* __restrict__ n_ex,
                                                      const unsigned int* __restrict__ ex_list, unsigned int ex_pitch,
                                                      const unsigned int* __restrict__ tag,
                                                      const unsigned int* __restrict__ rtag, unsigned int N,
                                                      unsigned int ex_start, const unsigned int* __restrict__ enable)
    {
    if (enable && !*enable) return;
    const unsigned int lane = threadIdx.x & 31u;
    const unsigned int sub = lane % G, seg = lane / G;
    const unsigned int i = (blockIdx.x * blockDim.x + threadIdx.x) / G;
    const bool live = i < N;
    unsigned int row = 0, nex = 0;
    if (live)
        {
        row = TAGKEYED ? tag[i] : i;
        nex = (TAGKEYED && row == 0xffffffffu) ? 0u : n_ex[row];
        }
    const bool work = live && ex_start < nex;
    if (!__any_sync(0xffffffffu, work)) return;
    unsigned int ex[FILTER_MAX_EX];
#pragma unroll
    for (int k = 0; k < FILTER_MAX_EX; k++)
        {
        ex[k] = 0xffffffffu;
        if (work && ex_start + k < nex)
            {
            const unsigned int e = ex_list[(size_t)(ex_start + k) * ex_pitch + row];
            ex[k] = TAGKEYED ? rtag[e] : e;
            }
        }
    const unsigned int head = work ? head_list[i] : 0u;
    const unsigned int n = work ? n_neigh[i] : 0u;
    unsigned int nmax = n;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) nmax = max(nmax, __shfl_xor_sync(0xffffffffu, nmax, o));
    const unsigned int segmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (seg * G));
    const unsigned int below = segmask & ((1u << lane) - 1u);
    unsigned int out = 0;
    for (unsigned int base = 0; base < nmax; base += G)
        {
        const unsigned int k = base + sub;
        unsigned int j = 0xffffffffu;
        bool keep = false;
        if (k < n)
            {
            j = nlist[head + k];
            keep = true;
#pragma unroll
            for (int e = 0; e < FILTER_MAX_EX; e++) keep &= (j != ex[e]);
            }
        const unsigned int mask = __ballot_sync(0xffffffffu, keep);
        // every lane has read its entry of this chunk before any lane writes (ballot = warp-wide sync point); writes land
        // at positions <= the chunk's own entries, which are not read again
        if (keep)
            {
            const unsigned int dst = out + __popc(mask & below);
            if (dst != k) nlist[head + dst] = j;
            }
        out += __popc(mask & segmask);
        }
    if (work && sub == 0 && out != n) n_neigh[i] = out;
    }

template<int G>
static void filter_launch(unsigned int* n_neigh, unsigned int* nlist, const unsigned int* head_list,
                          const unsigned int* n_ex, const unsigned int* ex_list, unsigned int ex_pitch, unsigned int start,
                          const unsigned int* tag, const unsigned int* rtag, unsigned int N, const unsigned int* enable,
                          cudaStream_t stream)
    {
    const unsigned int grid = (unsigned int)(((size_t)N * G + 255) / 256);
    if (tag)
        k_nlist_filter<true, G><<<grid, 256, 0, stream>>>(n_neigh, nlist, head_list, n_ex, ex_list, ex_pitch, tag, rtag, N,
                                                          start, enable);
    else
        k_nlist_filter<false, G><<<grid, 256, 0, stream>>>(n_neigh, nlist, head_list, n_ex, ex_list, ex_pitch, nullptr,
                                                           nullptr, N, start, enable);
    }

// row_hint: typical row length (Nmax); rows up to ~32 entries take 8 lanes each
cudaError_t launch_nlist_filter(unsigned int* n_neigh, unsigned int* nlist, const unsigned int* head_list,
                                const unsigned int* n_ex, const unsigned int* ex_list, unsigned int ex_pitch,
                                unsigned int max_ex, const unsigned int* tag, const unsigned int* rtag, unsigned int N,
                                const unsigned int* enable, cudaStream_t stream, unsigned int row_hint)
    {
    if (N == 0) return cudaSuccess;
    for (unsigned int start = 0; start < max_ex; start += FILTER_MAX_EX)
        {
        if (row_hint && row_hint <= 32)
            filter_launch<8>(n_neigh, nlist, head_list, n_ex, ex_list, ex_pitch, start, tag, rtag, N, enable, stream);
        else
            filter_launch<32>(n_neigh, nlist, head_list, n_ex, ex_list, ex_pitch, start, tag, rtag, N, enable, stream);
        }
    return cudaGetLastError();
    }

// ---- index-keyed copies of the tag-keyed tables ----------------------------------------------------------------------------
// What BondedGroupData::rebuildGPUTable / NeighborList::updateExListIdx do after a particle sort: row of particle i =
// row of tag[i] with every partner tag translated through rtag.  One launch per sort (every 300 steps) buys the per-step
// bond kernel and the per-rebuild filter a two-level load chain (row -> partner position) instead of four
// (tag -> row -> rtag -> position): k_bond_forces 36 -> see profiles/ at 1M beads.
__global__ void __launch_bounds__(256) k_rekey_tables(unsigned int* __restrict__ bond_n_idx, uint2* __restrict__ bond_list_idx,
                                                      const unsigned int* __restrict__ bond_n,
                                                      const uint2* __restrict__ bond_list, unsigned int bond_max,
                                                      unsigned int* __restrict__ ex_n_idx,
                                                      unsigned int* __restrict__ ex_list_idx,
                                                      const unsigned int* __restrict__ ex_n,
                                                      const unsigned int* __restrict__ ex_list, unsigned int ex_max,
                                                      unsigned int pitch, const unsigned int* __restrict__ tag,
                                                      const unsigned int* __restrict__ rtag, unsigned int N)
    {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const unsigned int t = tag[i];
    if (bond_n)
        {
        const unsigned int nb = bond_n[t];
        bond_n_idx[i] = nb;
        for (unsigned int k = 0; k < nb && k < bond_max; k++)
            {
            const uint2 e = bond_list[(size_t)k * pitch + t];
            bond_list_idx[(size_t)k * pitch + i] = make_uint2(rtag[e.x], e.y);
            }
        }
    if (ex_n)
        {
        const unsigned int ne = ex_n[t];
        ex_n_idx[i] = ne;
        for (unsigned int k = 0; k < ne && k < ex_max; k++)
            ex_list_idx[(size_t)k * pitch + i] = rtag[ex_list[(size_t)k * pitch + t]];
        }
    }

cudaError_t launch_rekey_tables(unsigned int* bond_n_idx, unsigned int* bond_list_idx, const unsigned int* bond_n,
                                const unsigned int* bond_list, unsigned int bond_max, unsigned int* ex_n_idx,
                                unsigned int* ex_list_idx, const unsigned int* ex_n, const unsigned int* ex_list,
                                unsigned int ex_max, unsigned int pitch, const unsigned int* tag, const unsigned int* rtag,
                                unsigned int N, cudaStream_t stream)
    {
    if (N == 0) return cudaSuccess;
    k_rekey_tables<<<(N + 255) / 256, 256, 0, stream>>>(bond_n_idx, (uint2*)bond_list_idx, bond_n, (const uint2*)bond_list,
                                                        bond_max, ex_n_idx, ex_list_idx, ex_n, ex_list, ex_max, pitch, tag,
                                                        rtag, N);
    return cudaGetLastError();
    }

// ---- net_force += force (second ForceCompute of the integrator) -------------------------------------------------------------
__global__ void __launch_bounds__(256) k_add_force(float4* __restrict__ net, float* __restrict__ net_virial,
                                                   unsigned int net_pitch, const float4* __restrict__ f,
                                                   const float* __restrict__ v, unsigned int v_pitch, unsigned int N,
                                                   bool virial)
    {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    float4 a = net[i];
    const float4 b = f[i];
    a.x = __fadd_rn(a.x, b.x); a.y = __fadd_rn(a.y, b.y); a.z = __fadd_rn(a.z, b.z); a.w = __fadd_rn(a.w, b.w);
    net[i] = a;
    if (virial)
        {
#pragma unroll
        for (int c = 0; c < 6; c++)
            net_virial[(size_t)c * net_pitch + i] = __fadd_rn(net_virial[(size_t)c * net_pitch + i], v[(size_t)c * v_pitch + i]);
        }
    }

cudaError_t launch_add_force(float4* net, float* net_virial, unsigned int net_pitch, const float4* f, const float* v,
                             unsigned int v_pitch, unsigned int N, bool virial, cudaStream_t stream)
    {
    if (N == 0) return cudaSuccess;
    k_add_force<<<(N + 255) / 256, 256, 0, stream>>>(net, net_virial, net_pitch, f, v, v_pitch, N, virial);
    return cudaGetLastError();
    }
} // namespace b200

extern "C" {
int b200_compute_fene_bond_forces(const b200_bond_args* args, const float* d_params, unsigned int* d_flags,
                                  b200_stream_t stream)
    {
    return (int)b200::bond_forces<b200::EvaluatorBondFENE>(args, d_params, d_flags, nullptr, (cudaStream_t)stream);
    }

int b200_compute_harmonic_bond_forces(const b200_bond_args* args, const float* d_params, unsigned int* d_flags,
                                      b200_stream_t stream)
    {
    return (int)b200::bond_forces<b200::EvaluatorBondHarmonic>(args, d_params, d_flags, nullptr, (cudaStream_t)stream);
    }

int b200_nlist_filter(unsigned int* d_n_neigh, unsigned int* d_nlist, const unsigned int* d_head_list,
                      const unsigned int* d_n_ex, const unsigned int* d_ex_list, unsigned int ex_pitch,
                      unsigned int ex_height, const unsigned int* d_tag, const unsigned int* d_rtag, unsigned int N,
                      b200_stream_t stream)
    {
    if (!d_n_neigh || !d_nlist || !d_head_list || !d_n_ex || !d_ex_list) return (int)cudaErrorInvalidValue;
    if ((d_tag == nullptr) != (d_rtag == nullptr)) return (int)cudaErrorInvalidValue;
    // batches of FILTER_MAX_EX exclusions per pass, like the reference driver (hoomd/md/NeighborListGPU.cu:226-240)
    return (int)b200::launch_nlist_filter(d_n_neigh, d_nlist, d_head_list, d_n_ex, d_ex_list, ex_pitch, ex_height, d_tag,
                                          d_rtag, N, nullptr, (cudaStream_t)stream);
    }
}
