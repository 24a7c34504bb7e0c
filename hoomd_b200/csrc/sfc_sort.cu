// SFCPackUpdater for sm_100a (replaces hoomd/SFCPackUpdaterGPU.cu:16-242).
// The reference builds a grid^3 table of Hilbert traversal indices on the host by recursion
// (hoomd/SFCPackUpdater.cc:245-441, 64 MB at grid 256) and looks each particle's bin up in it.  Here every thread walks
// the same recursion for its own bin (log2(grid) levels, the 8-entry cell order kept packed in one register), so there
// is no table and the key equals the reference's traversal index: the sorted order is IDENTICAL to the reference's
// (ties broken by particle index, as std::sort of (key, index) pairs does, hoomd/SFCPackUpdater.cc:571-583).
#include "common.cuh"
#include <cub/device/device_radix_sort.cuh>

namespace b200
{
// permute1..8 (hoomd/SFCPackUpdater.cc:263-366): result[r] = in[P[m][r]], 3 bits per entry
__device__ __constant__ unsigned char c_perm[8][8] = {
    {0, 3, 4, 7, 6, 5, 2, 1}, {0, 7, 6, 1, 2, 5, 4, 3}, {0, 7, 6, 1, 2, 5, 4, 3}, {2, 3, 0, 1, 6, 7, 4, 5},
    {2, 3, 0, 1, 6, 7, 4, 5}, {4, 3, 2, 5, 6, 1, 0, 7}, {4, 3, 2, 5, 6, 1, 0, 7}, {6, 5, 2, 1, 0, 3, 4, 7}};
// inverse of the (istep, jstep, kstep) walking tables (hoomd/SFCPackUpdater.cc:246-250): (i<<2 | j<<1 | k) -> sub-cell code
__device__ __constant__ unsigned char c_code_of_ijk[8] = {0, 1, 3, 2, 7, 6, 4, 5};

__device__ __forceinline__ unsigned int hilbert_traversal_index(unsigned int ib, unsigned int jb, unsigned int kb,
                                                                int levels)
    {
    unsigned int order = 0; // cell_order[m] in bits [3m, 3m+3), seed 0..7
#pragma unroll
    for (int m = 0; m < 8; m++) order |= (unsigned int)m << (3 * m);
    unsigned int index = 0;
    for (int l = levels - 1; l >= 0; l--)
        {
        const unsigned int code = c_code_of_ijk[(((ib >> l) & 1u) << 2) | (((jb >> l) & 1u) << 1) | ((kb >> l) & 1u)];
        unsigned int m = 0;
#pragma unroll
        for (int q = 0; q < 8; q++)
            if (((order >> (3 * q)) & 7u) == code) m = q;
        index = (index << 3) | m;
        unsigned int child = 0;
#pragma unroll
        for (int r = 0; r < 8; r++) child |= ((order >> (3 * c_perm[m][r])) & 7u) << (3 * r);
        order = child;
        }
    return index;
    }

__global__ void __launch_bounds__(256) k_sfc_keys(unsigned int* __restrict__ keys, unsigned int* __restrict__ idx,
                                                  const float4* __restrict__ pos, unsigned int N, Box box,
                                                  unsigned int grid, int levels)
    {
    unsigned int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const float4 p = pos[n];
    const float3 f = make_fraction_exact(make_float3(p.x, p.y, p.z), box, make_float3(0.0f, 0.0f, 0.0f));
    // hoomd/SFCPackUpdater.cc:551-565
    int ib = (int)((unsigned int)__fmul_rn(f.x, (float)grid) % grid);
    int jb = (int)((unsigned int)__fmul_rn(f.y, (float)grid) % grid);
    int kb = (int)((unsigned int)__fmul_rn(f.z, (float)grid) % grid);
    keys[n] = hilbert_traversal_index((unsigned int)ib, (unsigned int)jb, (unsigned int)kb, levels);
    idx[n] = n;
    }

struct float3s { float x, y, z; };
struct int3s { int x, y, z; };

__global__ void __launch_bounds__(256) k_sfc_apply(b200_pdata_arrays out, b200_pdata_arrays in,
                                                   unsigned int* __restrict__ rtag,
                                                   const unsigned int* __restrict__ order, unsigned int N)
    {
    unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const unsigned int o = order[i];
    if (in.pos) reinterpret_cast<float4*>(out.pos)[i] = reinterpret_cast<const float4*>(in.pos)[o];
    if (in.vel) reinterpret_cast<float4*>(out.vel)[i] = reinterpret_cast<const float4*>(in.vel)[o];
    if (in.accel) reinterpret_cast<float3s*>(out.accel)[i] = reinterpret_cast<const float3s*>(in.accel)[o];
    if (in.image) reinterpret_cast<int3s*>(out.image)[i] = reinterpret_cast<const int3s*>(in.image)[o];
    if (in.net_force) reinterpret_cast<float4*>(out.net_force)[i] = reinterpret_cast<const float4*>(in.net_force)[o];
    if (in.charge) out.charge[i] = in.charge[o];
    if (in.diameter) out.diameter[i] = in.diameter[o];
    if (in.body) out.body[i] = in.body[o];
    if (in.tag)
        {
        const unsigned int t = in.tag[o];
        out.tag[i] = t;
        if (rtag) rtag[t] = i;
        }
    }
} // namespace b200

using namespace b200;

extern "C" {

int b200_sfc_generate_sorted_order(unsigned int* d_sort_order, unsigned int* d_keys_scratch, unsigned int* d_scratch2x,
                                   void* d_temp, uint64_t temp_bytes, uint64_t* temp_bytes_needed, const float* d_pos,
                                   unsigned int N, const b200_box* box, unsigned int grid, b200_stream_t stream_)
    {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (grid == 0 || (grid & (grid - 1)) || grid > 1024) return (int)cudaErrorInvalidValue;
    int levels = 0;
    while ((1u << levels) < grid) levels++;
    // d_scratch2x: [0,N) keys out, [N,2N) identity indices in
    size_t need = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, need, d_keys_scratch, d_scratch2x, d_scratch2x + N, d_sort_order, (int)N, 0,
                                    3 * levels, stream);
    if (temp_bytes_needed) *temp_bytes_needed = need;
    if (!d_temp) return 0; // size query
    if (temp_bytes < need) return (int)cudaErrorInvalidValue;
    if (!box_is_orthorhombic(*box)) return (int)cudaErrorNotSupported;
    if (N == 0) return 0;
    k_sfc_keys<<<(N + 255) / 256, 256, 0, stream>>>(d_keys_scratch, d_scratch2x + N, (const float4*)d_pos, N,
                                                    make_box(*box), grid, levels);
    size_t tb = temp_bytes;
    cudaError_t err = cub::DeviceRadixSort::SortPairs(d_temp, tb, d_keys_scratch, d_scratch2x, d_scratch2x + N,
                                                      d_sort_order, (int)N, 0, 3 * levels, stream);
    if (err != cudaSuccess) return (int)err;
    return (int)cudaGetLastError();
    }

int b200_sfc_apply_sorted_order(const b200_pdata_arrays* out, const b200_pdata_arrays* in, unsigned int* d_rtag,
                                const unsigned int* d_sort_order, unsigned int N, b200_stream_t stream)
    {
    if (N == 0) return 0;
    k_sfc_apply<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(*out, *in, d_rtag, d_sort_order, N);
    return (int)cudaGetLastError();
    }

} // extern "C"
