// Device helpers shared by every kernel: box arithmetic in the reference's exact FP32 operation order, Saru RNG.
// Everything that decides MEMBERSHIP or is compared bit-for-bit against the CPU oracle uses the *_rn intrinsics so
// nvcc cannot contract a*b+c into an FMA (the oracle is compiled with -ffp-contract=off, see oracle/Makefile).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/hoomd_b200.h"

#define B200_NOT_BINNED 0xffffffffu

namespace b200
{
// ---- programmatic dependent launch (sm_90+) -------------------------------------------------------------------------------
// The rebuild bracket is a chain of small dependent kernels; with the attribute below a kernel's CTAs are scheduled while
// its predecessor is still running and block in pdl_wait() until that grid has completed and flushed its writes.  Every
// kernel of such a chain calls pdl_wait() before its first global access (completion is then transitive along the chain)
// and pdl_trigger() right after it.  Both instructions are no-ops in an ordinary launch.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template<class... KArgs, class... Args>
inline cudaError_t launch_chain(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                bool pdl, Args&&... args)
    {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
    }
#endif
bool pdl_enabled(); // HOOMD_B200_PDL (cell_list.cu)

// The box a kernel works in: `lo/hi/L/Linv/periodic` locate particles (binning, wrapping); the i-fields are the box MINIMUM
// IMAGES are taken in.  On one GPU both are the same box.  Under domain decomposition the first is the rank's local box
// (non-periodic along split axes) and the second the GLOBAL box: ghosts hold their owners' raw coordinates and a pair
// separation is pos_i - pos_j followed by the global minimum image -- the arithmetic of the single-domain reference
// (hoomd/md/PotentialPair.h:337 takes getGlobalBox() as well), so results do not depend on the decomposition.
// iforce (split axes): bit 0 / bit 1 = the local box's low / high face lies on the global periodic boundary, so the ghost
// cells on that side hold raw coordinates from the far side of the global box; ighost = how many cell layers from that face
// have such a cell among their neighbors (set by the neighbor-list launcher from the ghost width; 255 = all).
struct Box
    {
    float3 lo, hi, L, Linv;
    uchar3 periodic;
    float3 ilo, ihi, iL, iLinv;
    uchar3 iperiodic, iforce, ighost;
    };

__host__ __device__ inline Box make_box(const b200_box& b)
    {
    Box r;
    r.lo = make_float3(b.lo[0], b.lo[1], b.lo[2]);
    r.hi = make_float3(b.hi[0], b.hi[1], b.hi[2]);
    r.L = make_float3(b.L[0], b.L[1], b.L[2]);
    r.Linv = make_float3(b.Linv[0], b.Linv[1], b.Linv[2]);
    r.periodic = make_uchar3(b.periodic[0], b.periodic[1], b.periodic[2]);
    r.ilo = r.lo; r.ihi = r.hi; r.iL = r.L; r.iLinv = r.Linv;
    r.iperiodic = r.periodic;
    r.iforce = make_uchar3(0, 0, 0);
    r.ighost = make_uchar3(255, 255, 255);
    return r;
    }

// local box `b` for locating particles, minimum images in `img` (NULL: in `b` itself)
__host__ __device__ inline Box make_box_image(const b200_box& b, const b200_box* img)
    {
    Box r = make_box(b);
    if (img)
        {
        r.ilo = make_float3(img->lo[0], img->lo[1], img->lo[2]);
        r.ihi = make_float3(img->hi[0], img->hi[1], img->hi[2]);
        r.iL = make_float3(img->L[0], img->L[1], img->L[2]);
        r.iLinv = make_float3(img->Linv[0], img->Linv[1], img->Linv[2]);
        r.iperiodic = make_uchar3(img->periodic[0], img->periodic[1], img->periodic[2]);
        unsigned char f[3];
        for (int a = 0; a < 3; a++)
            f[a] = (img->periodic[a] && !b.periodic[a]) ? (unsigned char)((b.lo[a] <= img->lo[a] ? 1 : 0) | (b.hi[a] >= img->hi[a] ? 2 : 0)) : (unsigned char)0;
        r.iforce = make_uchar3(f[0], f[1], f[2]);
        }
    return r;
    }

inline bool box_is_orthorhombic(const b200_box& b) { return b.xy == 0.0f && b.xz == 0.0f && b.yz == 0.0f; }

// BoxDim::minImage, host form (hoomd/BoxDim.h:301-348) for an orthorhombic box and |w| < 1.5 L:
// z: compare-and-shift; x,y: `int i = (w*Linv+0.5); w -= i*L` which is i == 1 in that range => the same shift.
__device__ __forceinline__ float min_image_1d(float w, float lo, float hi, float L, float Linv, bool periodic_xy)
    {
    if (w >= hi)
        {
        if (periodic_xy)
            {
            int i = (int)__fadd_rn(__fmul_rn(w, Linv), 0.5f);
            w = __fsub_rn(w, __fmul_rn((float)i, L));
            }
        else
            w = __fsub_rn(w, L);
        }
    else if (w < lo)
        {
        if (periodic_xy)
            {
            int i = (int)__fadd_rn(__fmul_rn(-w, Linv), 0.5f);
            w = __fadd_rn(w, __fmul_rn((float)i, L));
            }
        else
            w = __fadd_rn(w, L);
        }
    return w;
    }

// Exact oracle form (used where membership is decided: neighbor list, distance check)
__device__ __forceinline__ float3 min_image_exact(float3 w, const Box& box)
    {
    if (box.iperiodic.z) w.z = min_image_1d(w.z, box.ilo.z, box.ihi.z, box.iL.z, box.iLinv.z, false);
    if (box.iperiodic.y) w.y = min_image_1d(w.y, box.ilo.y, box.ihi.y, box.iL.y, box.iLinv.y, true);
    if (box.iperiodic.x) w.x = min_image_1d(w.x, box.ilo.x, box.ihi.x, box.iL.x, box.iLinv.x, true);
    return w;
    }

// Branch-free restatement of min_image_exact for candidates whose separation satisfies |w| < 1.5 L (both particles inside
// the box): there `int i = (int)(|w|*Linv + 0.5f)` can only be 0 or 1, so `(float)i * L` is L when the rounded sum
// reaches 1.0f and 0 otherwise; w -+ 0 leaves w unchanged (up to the sign of zero, which dot() erases).  Same bits out.
__device__ __forceinline__ float min_image_1d_sel(float w, float lo, float hi, float L, float Linv, bool periodic_xy)
    {
    const bool up = w >= hi, dn = w < lo;
    float s = L;
    if (periodic_xy)
        {
        const float val = __fadd_rn(__fmul_rn(up ? w : -w, Linv), 0.5f);
        s = (val >= 1.0f) ? L : 0.0f;
        }
    return __fadd_rn(w, up ? -s : (dn ? s : 0.0f));
    }

__device__ __forceinline__ float3 min_image_exact_sel(float3 w, const Box& box, bool wx, bool wy, bool wz)
    {
    if (wz) w.z = min_image_1d_sel(w.z, box.ilo.z, box.ihi.z, box.iL.z, box.iLinv.z, false);
    if (wy) w.y = min_image_1d_sel(w.y, box.ilo.y, box.ihi.y, box.iL.y, box.iLinv.y, true);
    if (wx) w.x = min_image_1d_sel(w.x, box.ilo.x, box.ihi.x, box.iL.x, box.iLinv.x, true);
    return w;
    }

// which axes of a cell's candidates need the minimum image at all: cells on a periodic face of the image box, short axes,
// and -- under domain decomposition -- the cell layers of a split axis whose neighbors include ghost cells with raw
// coordinates from across the global boundary
__device__ __forceinline__ bool cell_needs_image(bool iperiodic, bool local_periodic, unsigned int iforce, unsigned int ighost,
                                                 unsigned int dim, unsigned int b)
    {
    if (!iperiodic) return false;
    if (local_periodic) return dim < 5 || b == 0 || b == dim - 1;
    // split axis: only next to a face that coincides with the global boundary
    return ((iforce & 1u) && b <= ighost) || ((iforce & 2u) && b + ighost + 1u >= dim);
    }

// Branch-free form for the force kernel: w -= L * rint(w / L), the reference's DEVICE minImage (hoomd/BoxDim.h:281-300).
// Identical to the host form whenever |w| != L/2 exactly (a pair at L/2 is beyond any legal cutoff: the box must be
// wider than 2*r_list, hoomd/md/NeighborListBinned.cc:109-115).
__device__ __forceinline__ float3 min_image_fast(float3 w, const Box& box)
    {
    if (box.iperiodic.z) w.z = __fmaf_rn(-box.iL.z, rintf(__fmul_rn(w.z, box.iLinv.z)), w.z);
    if (box.iperiodic.y) w.y = __fmaf_rn(-box.iL.y, rintf(__fmul_rn(w.y, box.iLinv.y)), w.y);
    if (box.iperiodic.x) w.x = __fmaf_rn(-box.iL.x, rintf(__fmul_rn(w.x, box.iLinv.x)), w.x);
    return w;
    }

// true when p is at least `margin` away from every periodic face of the box (margin <= 0: never)
__device__ __forceinline__ bool box_interior(float4 p, const Box& box, float margin)
    {
    if (!(margin > 0.0f)) return false;
    bool in = true;
    if (box.iperiodic.x) in &= (p.x >= box.ilo.x + margin) && (p.x < box.ihi.x - margin);
    if (box.iperiodic.y) in &= (p.y >= box.ilo.y + margin) && (p.y < box.ihi.y - margin);
    if (box.iperiodic.z) in &= (p.z >= box.ilo.z + margin) && (p.z < box.ihi.z - margin);
    return in;
    }

// dot(dx,dx) exactly as HOOMDMath.h:324-327 evaluates it on the host: (x*x + y*y) + z*z, three rounded products
__device__ __forceinline__ float dot_exact(float3 a)
    {
    return __fadd_rn(__fadd_rn(__fmul_rn(a.x, a.x), __fmul_rn(a.y, a.y)), __fmul_rn(a.z, a.z));
    }

// BoxDim::makeFraction (hoomd/BoxDim.h:238-245), orthorhombic: ((v - lo) * Linv + gf) / (1 + 2 gf), gf = ghost_width / L
__device__ __forceinline__ float3 make_fraction_exact(float3 v, const Box& box, float3 ghost_width)
    {
    float3 d = make_float3(__fsub_rn(v.x, box.lo.x), __fsub_rn(v.y, box.lo.y), __fsub_rn(v.z, box.lo.z));
    // delta.x -= (xz - yz*xy)*v.z + xy*v.y with zero tilts subtracts +-0: a no-op on every finite value
    float3 gf = make_float3(__fdiv_rn(ghost_width.x, box.L.x), __fdiv_rn(ghost_width.y, box.L.y),
                            __fdiv_rn(ghost_width.z, box.L.z));
    float3 f;
    f.x = __fdiv_rn(__fadd_rn(__fmul_rn(d.x, box.Linv.x), gf.x), __fadd_rn(1.0f, __fmul_rn(2.0f, gf.x)));
    f.y = __fdiv_rn(__fadd_rn(__fmul_rn(d.y, box.Linv.y), gf.y), __fadd_rn(1.0f, __fmul_rn(2.0f, gf.y)));
    f.z = __fdiv_rn(__fadd_rn(__fmul_rn(d.z, box.Linv.z), gf.z), __fadd_rn(1.0f, __fmul_rn(2.0f, gf.z)));
    return f;
    }

// Bin of a particle, following hoomd/CellList.cc:485-528 / hoomd/md/NeighborListBinned.cc:153-166.
// Returns false when the particle is outside the box + ghost layer (fraction outside [-1e-5, 1.00001) or bin out of range).
__device__ __forceinline__ bool particle_bin(float3 p, const Box& box, float3 ghost_width, uint3 dim, int3& bin)
    {
    float3 f = make_fraction_exact(p, box, ghost_width);
    int ib = (int)__fmul_rn(f.x, (float)dim.x);
    int jb = (int)__fmul_rn(f.y, (float)dim.y);
    int kb = (int)__fmul_rn(f.z, (float)dim.z);
    if ((f.x < -0.00001f || f.x >= 1.00001f) || (f.y < -0.00001f || f.y >= 1.00001f)
        || (f.z < -0.00001f || f.z >= 1.00001f))
        return false;
    if (ib == (int)dim.x && box.periodic.x) ib = 0;
    if (jb == (int)dim.y && box.periodic.y) jb = 0;
    if (kb == (int)dim.z && box.periodic.z) kb = 0;
    bin = make_int3(ib, jb, kb);
    if (ib < 0 || ib >= (int)dim.x || jb < 0 || jb >= (int)dim.y || kb < 0 || kb >= (int)dim.z) return false;
    return true;
    }

// BoxDim::wrap (hoomd/BoxDim.h:366-432), orthorhombic, flags = 0
__device__ __forceinline__ void box_wrap(float& x, float& y, float& z, int3& img, const Box& box)
    {
    if (box.periodic.x)
        {
        if (x >= box.hi.x) { x = __fsub_rn(x, box.L.x); img.x++; }
        else if (x < box.lo.x) { x = __fadd_rn(x, box.L.x); img.x--; }
        }
    if (box.periodic.y)
        {
        if (y >= box.hi.y) { y = __fsub_rn(y, box.L.y); img.y++; }
        else if (y < box.lo.y) { y = __fadd_rn(y, box.L.y); img.y--; }
        }
    if (box.periodic.z)
        {
        if (z >= box.hi.z) { z = __fsub_rn(z, box.L.z); img.z++; }
        else if (z < box.lo.z) { z = __fadd_rn(z, box.L.z); img.z--; }
        }
    }

// hoomd::detail::Saru (hoomd/Saru.h): 3-seed constructor :324-340, u32<1>() :417-424, f() :436-441, s(a,b) :476-481,580
struct Saru
    {
    unsigned int x, y;
    __host__ __device__ inline Saru(unsigned int seed1, unsigned int seed2, unsigned int seed3)
        {
        seed3 ^= (seed1 << 7) ^ (seed2 >> 6);
        seed2 += (seed1 >> 4) ^ (seed3 >> 15);
        seed1 ^= (seed2 << 9) + (seed3 << 8);
        seed3 ^= 0xA5366B4Du * ((seed2 >> 11) ^ (seed1 << 1));
        seed2 += 0x72BE1579u * ((seed1 << 4) ^ (seed3 >> 16));
        seed1 ^= 0x3F38A6EDu * ((seed3 >> 5) ^ (unsigned int)(((int)seed2) >> 22));
        seed2 += seed1 * seed3;
        seed1 += seed3 ^ (seed2 >> 2);
        seed2 ^= (unsigned int)(((int)seed2) >> 17);

        x = 0x79dedea3u * (seed1 ^ (unsigned int)(((int)seed1) >> 14));
        y = (x + seed2) ^ (unsigned int)(((int)x) >> 8);
        x = x + (y * (y ^ 0xdddf97f5u));
        y = 0xABCB96F7u + (y >> 1);
        }
    __host__ __device__ inline unsigned int u32()
        {
        x = 0x4beb5d59u * x + 0x2600e1f7u;                                   // advanceLCG<1>
        y = y + 0x8009d14bu + ((unsigned int)(((int)y) >> 31) & 0xda879addu); // advanceWeyl<1>
        const unsigned int v = (x ^ (x >> 26)) + y;
        return (v ^ (v >> 20)) * 0x6957f5a7u;
        }
    __device__ inline float f()
        {
        // (int)u32 * 2^-32 is exact (power-of-two scale), so the single rounding is the + 0.5f: FMA-safe
        return __fadd_rn(__fmul_rn((float)(int)u32(), 2.32830643653869628906250e-10f), 0.5f);
        }
    __device__ inline float s(float a, float b) { return __fadd_rn(a, __fmul_rn(__fsub_rn(b, a), f())); }
    };

#define B200_CHECK_LAUNCH() cudaGetLastError()

} // namespace b200
