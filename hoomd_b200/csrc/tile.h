// Cell-tile path (tile.cu): argument blocks and launchers shared with the engine.  Not part of the C ABI.
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace b200
{
constexpr int TILE_BLOCK = 128; // 4 warps x 32 owned particles per round

// geometry of the cell tiles: a CTA owns b.x * b.y * b.z cells and stages them plus one halo layer
struct TileGeom
    {
    uint3 dim;        // cell grid
    uint3 b;          // cells per tile along each axis; (b.x+2)(b.y+2)(b.z+2) <= 128, b.y*b.z <= 16
    uint3 tdim;       // tiles per axis = ceil(dim / b)
    unsigned int cap; // staged positions per CTA (shared-memory capacity)
    };

struct TileSortArgs
    {
    const unsigned int* enable;
    unsigned int N;
    const unsigned int* perm;   // CellList cell_idx: previous index of the particle now at e
    const float4* sorted_pos;   // CellList cell_xyzf with the type in w
    float4* pos; float4* last_pos;
    float4* vel; float4* vel_alt;
    float* accel; float* accel_alt;
    int* image; int* image_alt;
    unsigned int* tag; unsigned int* tag_alt;
    unsigned int* rtag;
    };

struct TileBuildArgs
    {
    const unsigned int* enable;
    TileGeom geom;
    const unsigned int* cell_start;
    const float4* pos;          // cell-sorted positions (type in w)
    uint4* ell4;                // chunk c (8 slots, 16 B) of particle g: ell4[c * pitch + g]; last chunk padded with slot geom.cap
    unsigned int pitch, Nmax;   // Nmax is a multiple of 8
    unsigned int* nn;           // neighbors per particle (<= Nmax)
    unsigned int* conditions;   // [0] longest row seen, [1] largest halo population that exceeded geom.cap
    const float* r_cut; const float* r_listsq; unsigned int ntypes;
    Box box;
    };

struct TileStepArgs
    {
    TileGeom geom;
    const unsigned int* cell_start;
    const uint4* ell4;
    unsigned int pitch, tpp, Nmax;
    const unsigned int* nn;
    int force_only;
    };

cudaError_t launch_tile_sort(const TileSortArgs& a, cudaStream_t stream);
cudaError_t launch_tile_build(const TileBuildArgs& a, cudaStream_t stream);
cudaError_t launch_tile_post(StepState* st, const unsigned int* conditions, unsigned int Nmax, cudaStream_t stream);
cudaError_t launch_tile_to_csr(const TileBuildArgs& a, unsigned int* out_n, unsigned int* out_head,
                               unsigned int* out_list, cudaStream_t stream);
cudaError_t launch_md_step_tile_lj(const b200_pair_args& pa, const float* d_params, const FusedStepArgs& fa,
                                   const TileStepArgs& ta, const float4* pos_in, float4* pos_out, int copy_pos_on_last,
                                   StepState* st, unsigned long long cond_handle, int use_cond, cudaStream_t stream);
} // namespace b200
