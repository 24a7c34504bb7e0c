/* hoomd_b200.h -- C ABI of the B200-native HOOMD-blue MD inner loop.
 *
 * Two layers, both plain C (pointers, sizes, POD structs; no C++/torch types):
 *
 *  1. KERNEL DRIVERS (b200_*): one entry point per reference GPU driver `gpu_*()`.  The reference's `...GPU.cc`
 *     host classes call these with raw device pointers taken from ArrayHandle<T>(..., access_location::device).
 *     Same convention as the reference drivers: return cudaError_t-compatible int (0 = cudaSuccess), never throw,
 *     never synchronise, caller owns every buffer, everything is enqueued on the given stream (0 = legacy default
 *     stream, which is what the reference uses everywhere).
 *
 *  2. ENGINE (b200_sys_*): a device-resident system (ParticleData + CellList + NeighborListBinned +
 *     PotentialPairLJ + IntegratorTwoStep{TwoStepNVE|TwoStepLangevin} + SFCPackUpdater + ComputeThermo) with the
 *     reference's step semantics (hoomd/System.cc:484-599, hoomd/md/IntegratorTwoStep.cc:96-161), driven without a
 *     host round trip per step (rebuild decision consumed on the device, whole step captured in a CUDA graph).
 *     This is what the python job-script surface (hoomd_b200.md.nlist.cell / md.pair.lj / md.integrate.*) drives.
 *
 * All floating point is FP32 (the reference's SINGLE_PRECISION build), indices/tags/counts are uint32, images int32x3.
 * Citations are relative to the reference tree (HOOMD-blue 2.3.0).
 */
#ifndef HOOMD_B200_H_
#define HOOMD_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_API __attribute__((visibility("default")))

typedef void* b200_stream_t; /* cudaStream_t */

/* ---------------------------------------------------------------------------------------------------------------
 * POD mirrors of the reference's by-value kernel arguments
 * ------------------------------------------------------------------------------------------------------------- */

/* BoxDim (hoomd/BoxDim.h:568-576).  Orthorhombic only in this round: xy = xz = yz = 0 (every BASELINE config);
 * a non-zero tilt makes the drivers return cudaErrorNotSupported (801). */
typedef struct b200_box
    {
    float lo[3];
    float hi[3];
    float L[3];
    float Linv[3];
    float xy, xz, yz;
    unsigned char periodic[3];
    unsigned char _pad;
    } b200_box;

/* BoxDim(Lx,Ly,Lz): hi = L/2, lo = -hi, Linv = 1/L in FP32, periodic = (1,1,1)  (hoomd/BoxDim.h:93-98,165-171) */
B200_API void b200_box_init(b200_box* box, float Lx, float Ly, float Lz);
/* BoxDim(lo,hi,periodic)  (hoomd/BoxDim.h:129-134,193-199) */
B200_API void b200_box_init_lohi(b200_box* box, const float lo[3], const float hi[3], const unsigned char periodic[3]);

/* Index3D(w,h,d): (i,j,k) -> (k*h + j)*w + i  (hoomd/Index1D.h) */
typedef struct b200_dim3 { unsigned int x, y, z; } b200_dim3;

/* ---------------------------------------------------------------------------------------------------------------
 * CellList  -- replaces gpu_compute_cell_list (hoomd/CellListGPU.cuh:21-41, kernel hoomd/CellListGPU.cu:40-156)
 * and gpu_sort_cell_list (hoomd/CellListGPU.cuh:43-56).
 *
 * B200 layout: cells are CONTIGUOUS RANGES of a cell-sorted particle array (count -> scan -> scatter), no Nmax
 * padding and therefore no overflow/retry loop (hoomd/CellList.cc:181-196).  Inside a cell, entries are ordered by
 * particle index, i.e. the order the CPU reference produces (hoomd/CellList.cc:474-563) and the order
 * nlist.cell(deterministic=True) restores on the GPU (hoomd/CellListGPU.cc:92-118).
 *
 *   cell_start[ncells+1]   exclusive prefix sum of the cell occupancies (cell c = [cell_start[c], cell_start[c+1]))
 *   cell_xyzf [N+n_ghost]  (x, y, z, flag) per cell-sorted entry; flag = particle index bits (default), type or charge
 *   cell_type [N+n_ghost]  type of each entry (the .x of the reference's tdb array), optional
 *   cell_idx  [N+n_ghost]  particle index of each entry, optional (reference: getIndexArray)
 *
 * conditions (uint3, same meaning as hoomd/CellList.cc:575-621): .x = largest cell occupancy seen,
 * .y = index+1 of a particle with a NaN coordinate, .z = index+1 of a local particle outside the box.
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct b200_cell_list_args
    {
    /* outputs */
    unsigned int* d_cell_start;   /* [ncells + 1] */
    float*        d_cell_xyzf;    /* [4 * (N + n_ghost)] */
    unsigned int* d_cell_type;    /* [N + n_ghost] or NULL */
    unsigned int* d_cell_idx;     /* [N + n_ghost] or NULL */
    unsigned int* d_conditions;   /* [3] ; caller zeroes it (resetConditions) */
    /* scratch owned by the caller */
    unsigned int* d_cell_count;   /* [ncells] */
    unsigned int* d_particle_cell;/* [N + n_ghost] cell of every particle (0xffffffff = not binned) */
    unsigned int* d_particle_slot;/* [N + n_ghost] */
    unsigned int* d_scratch_idx;  /* [N + n_ghost] */
    /* inputs */
    const float*  d_pos;          /* Scalar4[N + n_ghost] */
    const float*  d_charge;       /* Scalar[N + n_ghost] or NULL (needed when flag_charge) */
    unsigned int  N;
    unsigned int  n_ghost;
    int           flag_charge;    /* CellList::setFlagCharge */
    int           flag_type;      /* CellList::setFlagType */
    b200_box      box;
    b200_dim3     dim;            /* CellList::getDim() */
    float         ghost_width[3]; /* CellList::getGhostWidth() */
    /* domain decomposition with RAW ghost coordinates (see b200_nlist_args.image_box): ghost g (particle N + g) is binned at
     * pos + d_ghost_shift[3 * d_ghost_dir[g] ...] -- the periodic image that lies next to the local box -- while the stored
     * entry keeps the raw position.  Both NULL: ghosts are binned where they are (the reference's pre-wrapped ghosts). */
    const unsigned char* d_ghost_dir;   /* [n_ghost] or NULL */
    const float*         d_ghost_shift; /* [3 * 27] or NULL */
    /* NULL, or two words {0, 0} owned by the caller (zeroed ONCE, then left alone): with HOOMD_B200_CELL_FUSED=1 small systems
     * (<= 262,144 entries) are built by ONE cooperative launch -- bin, scan, scatter, rank separated by a grid-wide barrier on
     * these words -- instead of four launches; same result bit for bit, measured slower, so opt-in.
     * Builds that may overlap on the device must not share the words. */
    unsigned int*        d_fused_barrier;
    } b200_cell_list_args;

B200_API int b200_compute_cell_list(const b200_cell_list_args* args, b200_stream_t stream);

/* CellList::computeDimensions (hoomd/CellList.cc:69-131), host arithmetic only.  n_extra_ghost_cells: whether the box
 * is a domain-decomposed local box whose non-periodic directions get extra ghost cells (hoomd/CellList.cc:103-118). */
B200_API void b200_cell_list_dimensions(const b200_box* box, float nominal_width, unsigned int multiple,
                                        float ghost_layer_width, b200_dim3* dim, float ghost_width[3]);

/* Reference-layout export: cell_size[ncells], xyzf[Nmax*ncells] (cli(offset,cell) = cell*Nmax + offset), and the
 * (type, diameter, body, 0) tdb array, exactly as getCellSizeArray()/getXYZFArray()/getTDBArray() hold them
 * (hoomd/CellList.h:263-300).  Entries beyond Nmax are dropped and conditions.x records the needed Nmax, like the
 * reference's overflow protocol. */
B200_API int b200_cell_list_export_padded(unsigned int* d_cell_size, float* d_xyzf, float* d_tdb,
                                          unsigned int* d_conditions, const unsigned int* d_cell_start,
                                          const float* d_cell_xyzf, const unsigned int* d_cell_idx,
                                          const float* d_pos, const float* d_diameter, const unsigned int* d_body,
                                          unsigned int ncells, unsigned int Nmax, b200_stream_t stream);

/* CellList::initializeCellAdj (hoomd/CellList.cc:373-436) on the device: per cell the sorted, de-duplicated list of
 * the (2r+1)^3 wrapped neighbor cells; n_adj = prod(min(dim, 2r+1)); layout cadji(a, cell) = cell*n_adj + a. */
B200_API int b200_cell_list_build_adjacency(unsigned int* d_cell_adj, b200_dim3 dim, unsigned int radius,
                                            unsigned int n_adj, b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------------------------
 * NeighborList  -- replaces gpu_nlist_build_head_list (hoomd/md/NeighborListGPU.cu:400-433),
 * gpu_compute_nlist_binned (hoomd/md/NeighborListGPUBinned.cu:570-628), gpu_nlist_needs_update_check_new
 * (hoomd/md/NeighborListGPU.cu:84-113).  Results follow the CPU oracle hoomd/md/NeighborListBinned.cc:79-240 bit for
 * bit: dx = pos_i - pos_j, host-form minImage (hoomd/BoxDim.h:301-348), dr^2 = (dx*dx + dy*dy) + dz*dz without FMA
 * contraction, inclusion iff dr^2 <= r_listsq(type_i,type_j), full storage mode, and the same ENTRY ORDER as the CPU
 * loop (neighbor cells in ascending cell id, occupants in ascending particle index).
 * ------------------------------------------------------------------------------------------------------------- */

/* head[i] = exclusive scan of Nmax[type_i]; *d_req_size = total (NeighborList::buildHeadList, NeighborList.cc:1336-1357) */
B200_API int b200_nlist_build_head_list(unsigned int* d_head_list, unsigned int* d_req_size_nlist,
                                        const unsigned int* d_Nmax, const float* d_pos, unsigned int N,
                                        unsigned int ntypes, b200_stream_t stream);

typedef struct b200_nlist_args
    {
    /* outputs */
    unsigned int* d_nlist;        /* [>= head[N-1] + Nmax[type_{N-1}]] */
    unsigned int* d_n_neigh;      /* [N] */
    float*        d_last_pos;     /* Scalar4[N] or NULL: (x,y,z,0) at build time (setLastUpdatedPos, NeighborList.cc:1044-1066) */
    unsigned int* d_conditions;   /* [ntypes]: max(n_neigh) per type where it exceeded Nmax[type] (NeighborListBinned.cc:225) */
    /* inputs */
    const unsigned int* d_Nmax;      /* [ntypes] */
    const unsigned int* d_head_list; /* [N] */
    const float*  d_pos;             /* Scalar4[N + n_ghost] */
    unsigned int  N;                 /* local particles (lists are built for these only) */
    const unsigned int* d_cell_start;
    const float*  d_cell_xyzf;       /* flag must be the particle index */
    const unsigned int* d_cell_type; /* may be NULL when ntypes == 1 */
    b200_dim3     dim;
    b200_box      box;
    float         ghost_width[3];
    const float*  d_r_cut;           /* [ntypes*ntypes] r_cut(i,j); <= 0 disables the pair (NeighborListBinned.cc:193) */
    const float*  d_r_listsq;        /* [ntypes*ntypes] (r_cut+r_buff)^2 computed like NeighborList::updateRList (NeighborList.cc:354-400) */
    unsigned int  ntypes;
    int           half_storage;      /* 1: keep only i<j (NeighborList::half); GPU classes always use 0 */
    unsigned int  cand_hint;         /* tuning only: expected particles in a cell's <=27-cell neighborhood (sizes the
                                        shared-memory staging of a pass; 0 = 512).  Larger neighborhoods take more passes. */
    const b200_box* image_box;       /* NULL: minimum images in `box`.  Domain decomposition: the GLOBAL box -- `box` (the local
                                        box, non-periodic along split axes) only locates cells, separations are pos_i - pos_j
                                        followed by the global minimum image, as the single-domain reference forms them; ghosts
                                        then carry their owners' raw coordinates and the list does not depend on the
                                        decomposition.  (The reference pre-wraps ghosts instead, hoomd/Communicator.cc
                                        exchangeGhosts + getShiftedBox, which rounds their coordinates once more.) */
    } b200_nlist_args;

B200_API int b200_compute_nlist_binned(const b200_nlist_args* args, b200_stream_t stream);

/* *d_result = checkn if any particle moved >= its threshold since d_last_pos (NeighborList::distanceCheck,
 * NeighborList.cc:964-1040): |minImage(pos - lambda*last_pos)|^2 >= max(0,((rcut_max_i+r_buff)*lambda_min - rcut_max_i)/2)^2 */
B200_API int b200_nlist_needs_update_check(unsigned int* d_result, const float* d_last_pos, const float* d_pos,
                                           unsigned int N, const b200_box* box, const float* d_rcut_max, float r_buff,
                                           unsigned int ntypes, float lambda_min, const float lambda[3],
                                           unsigned int checkn, b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------------------------
 * PotentialPair<EvaluatorPairLJ>  -- replaces gpu_compute_ljtemp_forces(const pair_args_t&, const Scalar2*)
 * (hoomd/md/AllDriverPotentialPairGPU.cuh:20-22, hoomd/md/PotentialPairGPU.cuh:26-98,579-689); arithmetic of
 * hoomd/md/PotentialPair.h:309-522 + hoomd/md/EvaluatorPairLJ.h:137-158.
 * ------------------------------------------------------------------------------------------------------------- */
enum { B200_SHIFT_NONE = 0, B200_SHIFT_SHIFT = 1, B200_SHIFT_XPLOR = 2 }; /* PotentialPair::energyShiftMode */

typedef struct b200_pair_args
    {
    float*        d_force;        /* Scalar4[N]: (fx, fy, fz, pe_i = sum_j V_ij/2) */
    float*        d_virial;       /* [6*virial_pitch]: xx,xy,xz,yy,yz,zz at [k*pitch + i]; may be NULL if !compute_virial */
    unsigned int  virial_pitch;
    unsigned int  N;
    const float*  d_pos;          /* Scalar4[N + n_ghost] */
    b200_box      box;
    const unsigned int* d_n_neigh;
    const unsigned int* d_nlist;
    const unsigned int* d_head_list;
    const float*  d_rcutsq;       /* [ntypes*ntypes] */
    const float*  d_ronsq;        /* [ntypes*ntypes] (xplor only) */
    unsigned int  ntypes;
    unsigned int  shift_mode;
    unsigned int  compute_virial;
    unsigned int  threads_per_particle; /* 0 = pick (1,2,4,8,16,32 accepted; the reference autotunes this) */
    unsigned int  block_size;           /* 0 = pick */
    float         wrap_skip_margin;     /* > 0: particles farther than this from every periodic face skip minImage
                                           (identity there).  Must be >= the largest |r_i - r_j| of any listed pair,
                                           i.e. max r_cut + 2 r_buff; 0 = always apply minImage */
    } b200_pair_args;

/* d_params: Scalar2[ntypes*ntypes] = (lj1, lj2) = (4 eps sigma^12, 4 alpha eps sigma^6) (hoomd/md/pair.py:602-609) */
B200_API int b200_compute_lj_forces(const b200_pair_args* args, const float* d_params, b200_stream_t stream);

/* The other PotentialPair<evaluator> instantiations of hoomd/md/AllPairPotentials.h:45-79 through the same kernel
 * (replaces gpu_compute_gauss_forces / _yukawa_ / _mie_ / _force_shifted_lj_ / _morse_forces,
 * hoomd/md/AllDriverPotentialPairGPU.cuh).  d_params is the reference's param_type array [ntypes*ntypes]:
 *   LJ (lj1, lj2) Scalar2 | GAUSS (epsilon, sigma) Scalar2 | YUKAWA (epsilon, kappa) Scalar2 |
 *   MIE (mie1, mie2, n, m) Scalar4 | FORCE_SHIFTED_LJ (lj1, lj2) Scalar2 | MORSE (D0, alpha, r0, -) Scalar4
 *   DPD (A, gamma) Scalar2: the conservative force of md.pair.dpd_conservative (EvaluatorPairDPDThermo.h:139-160)
 * (coefficient processing: hoomd/md/pair.py gauss :611-690, yukawa :835-900, mie :2188-2265, force_shifted_lj :1781-1850,
 * morse :1193-1260).  Energy shift modes as for LJ. */
enum { B200_PAIR_LJ = 0, B200_PAIR_GAUSS = 1, B200_PAIR_YUKAWA = 2, B200_PAIR_MIE = 3, B200_PAIR_FORCE_SHIFTED_LJ = 4,
       B200_PAIR_MORSE = 5, B200_PAIR_DPD = 6 };
B200_API int b200_compute_pair_forces(int evaluator, const b200_pair_args* args, const void* d_params,
                                      b200_stream_t stream);

/* PotentialPairDPDThermoDPD -- md.pair.dpd, the pairwise DPD thermostat -- replaces gpu_compute_dpdthermodpd_forces
 * (hoomd/md/AllDriverPotentialPairGPU.cuh, dpd_pair_args_t of hoomd/md/PotentialPairDPDThermoGPU.cuh; CPU loop
 * hoomd/md/PotentialPairDPDThermo.h:116-270, evaluator hoomd/md/EvaluatorPairDPDThermo.h:161-230).  Same operands as the pair
 * force plus what dpd_pair_args_t adds: neighbor velocities and tags, the user seed, the time step, kT and deltaT.  d_params =
 * Scalar2 (A, gamma) per type pair.  The random number of a pair comes from Saru(min tag, max tag, seed + timestep), so both
 * rows of a full neighbor list get the same draw; energy and virial are those of the conservative part.  No energy shift. */
typedef struct b200_dpd_thermo_args
    {
    const float*        d_vel;      /* Scalar4[N + n_ghost] */
    const unsigned int* d_tag;      /* [N + n_ghost] */
    unsigned int        seed;       /* m_seed: the user seed AFTER PotentialPairDPDThermo::setSeed's hash
                                       (PotentialPairDPDThermo.h:93-105; same hash as b200_langevin_hash_seed) */
    unsigned int        timestep;
    float               T;          /* m_T->getValue(timestep) */
    float               deltaT;
    } b200_dpd_thermo_args;
B200_API int b200_compute_dpd_thermo_forces(const b200_pair_args* args, const float* d_params,
                                            const b200_dpd_thermo_args* thermo, b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------------------------
 * NeighborList exclusions  -- replaces gpu_nlist_filter (hoomd/md/NeighborListGPU.cu:140-240); semantics of
 * NeighborList::filterNlist (hoomd/md/NeighborList.cc:1275-1328): every row is rewritten in place without its excluded
 * partners, order of the survivors kept, n_neigh updated.  Exclusion table: d_ex_list[k * ex_pitch + row], k < d_n_ex[row]
 * (Index2D exli(row, k), hoomd/Index1D.h).  d_tag == d_rtag == NULL: rows and entries are particle INDICES (the
 * reference's m_ex_list_idx, rebuilt by updateExListIdx after every sort).  Both non-NULL: rows are keyed by tag[i] and
 * entries are TAGS resolved through rtag on the fly (the reference's m_ex_list_tag used directly: no table rebuild).
 * ------------------------------------------------------------------------------------------------------------- */
B200_API int b200_nlist_filter(unsigned int* d_n_neigh, unsigned int* d_nlist, const unsigned int* d_head_list,
                               const unsigned int* d_n_ex, const unsigned int* d_ex_list, unsigned int ex_pitch,
                               unsigned int ex_height /* Index2D::getH(): longest exclusion row */,
                               const unsigned int* d_tag, const unsigned int* d_rtag, unsigned int N,
                               b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------------------------
 * PotentialBond<EvaluatorBondFENE | EvaluatorBondHarmonic>  -- replaces gpu_compute_fene_forces /
 * gpu_compute_harmonic_forces(const bond_args_t&, params, d_flags) (hoomd/md/AllDriverPotentialBondGPU.cuh:17-24,
 * hoomd/md/PotentialBondGPU.cuh:26-98,112-330); arithmetic of hoomd/md/PotentialBond.h:155-300 +
 * hoomd/md/EvaluatorBondFENE.h:95-133 / EvaluatorBondHarmonic.h:75-92.  One thread per particle walks its bond rows:
 * force_i = sum dx * force_divr, pe_i = sum bond_eng/2, virial_i = sum dx dx force_divr/2, dx = minImage(pos_i - pos_j).
 * *d_flags is set to 1 when a bond cannot be evaluated (FENE: r >= r_0, "bond out of bounds").
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct b200_bond_args
    {
    float*        d_force;        /* Scalar4[N] */
    float*        d_virial;       /* [6*virial_pitch]; may be NULL if !compute_virial */
    unsigned int  virial_pitch;
    unsigned int  N;
    const float*  d_pos;          /* Scalar4 */
    const float*  d_diameter;     /* [n] or NULL = every diameter 1.0 (ParticleData default) */
    b200_box      box;
    const unsigned int* d_bond_list; /* uint2 (partner, bond type) at [k * pitch + row] (group_storage<2>, gpu_table_indexer) */
    unsigned int  pitch;
    const unsigned int* d_n_bonds;   /* [rows] */
    unsigned int  n_bond_types;
    unsigned int  compute_virial;
    const unsigned int* d_tag;    /* NULL: rows/partners are particle indices (BondedGroupData GPU table);           */
    const unsigned int* d_rtag;   /* both set: rows keyed by tag[i], partners are tags looked up in rtag             */
    } b200_bond_args;

/* d_params: Scalar4[n_bond_types] = (K, r_0, lj1, lj2), lj1 = 4 eps sigma^12, lj2 = 4 eps sigma^6 (hoomd/md/bond.py:367-372) */
B200_API int b200_compute_fene_bond_forces(const b200_bond_args* args, const float* d_params, unsigned int* d_flags,
                                           b200_stream_t stream);
/* d_params: Scalar2[n_bond_types] = (K, r_0) (hoomd/md/bond.py:290-296) */
B200_API int b200_compute_harmonic_bond_forces(const b200_bond_args* args, const float* d_params,
                                               unsigned int* d_flags, b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Integrator  -- replaces gpu_integrator_sum_net_force (hoomd/Integrator.cu:115-153), gpu_nve_step_one/two
 * (hoomd/md/TwoStepNVEGPU.cu:120-150,392-430), gpu_langevin_step_two (hoomd/md/TwoStepLangevinGPU.cu:444-500).
 * Arithmetic follows the CPU oracle (hoomd/md/TwoStepNVE.cc:77-122,255-300, hoomd/md/TwoStepLangevin.cc:86-125,
 * 240-330) operation by operation with FMA contraction disabled, so one half step is bit-identical to the reference.
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct b200_force_list /* gpu_force_list, hoomd/Integrator.cuh:22-58 (torques: out of scope, point particles) */
    {
    const float* f[6];            /* Scalar4 force arrays, NULL = unused */
    const float* v[6];            /* virial arrays */
    unsigned int vpitch[6];
    } b200_force_list;

B200_API int b200_integrator_sum_net_force(float* d_net_force, float* d_net_virial, unsigned int net_virial_pitch,
                                           const b200_force_list* force_list, unsigned int nparticles, int clear,
                                           int compute_virial, b200_stream_t stream);

B200_API int b200_nve_step_one(float* d_pos, float* d_vel, const float* d_accel, int* d_image,
                               const unsigned int* d_group_members, unsigned int group_size, const b200_box* box,
                               float deltaT, int limit, float limit_val, int zero_force, b200_stream_t stream);

B200_API int b200_nve_step_two(float* d_vel, float* d_accel, const unsigned int* d_group_members,
                               unsigned int group_size, const float* d_net_force, float deltaT, int limit,
                               float limit_val, int zero_force, b200_stream_t stream);

typedef struct b200_langevin_args /* langevin_step_two_args, hoomd/md/TwoStepLangevinGPU.cuh:18-34 */
    {
    const float*  d_gamma;        /* [n_types] */
    unsigned int  n_types;
    int           use_lambda;
    float         lambda;
    float         T;              /* kT at this timestep */
    unsigned int  timestep;
    unsigned int  seed;           /* ALREADY hashed as in TwoStepLangevinBase ctor (TwoStepLangevinBase.cc:54) */
    float*        d_sum_bdenergy; /* [1] or NULL; reservoir energy transfer of this step when tally */
    int           noiseless_t;
    int           tally;
    } b200_langevin_args;

/* The user-seed hash of the TwoStepLangevinBase constructor (hoomd/md/TwoStepLangevinBase.cc:54) */
B200_API unsigned int b200_langevin_hash_seed(unsigned int user_seed);

B200_API int b200_langevin_step_two(const float* d_pos, float* d_vel, float* d_accel, const float* d_diameter,
                                    const unsigned int* d_tag, const unsigned int* d_group_members,
                                    unsigned int group_size, const float* d_net_force,
                                    const b200_langevin_args* args, float deltaT, unsigned int D,
                                    b200_stream_t stream);

/* -----------------------------------------------------------------------------------------------------------------------
 * TwoStepNVTMTK (Nose-Hoover / MTK thermostat, point particles) -- replaces gpu_nvt_mtk_step_one / gpu_nvt_mtk_step_two
 * (hoomd/md/TwoStepNVTMTKGPU.cuh:17-38, .cu:30-212) and the host-side advanceThermostat (hoomd/md/TwoStepNVTMTK.cc:387-404).
 * The reference passes m_exp_thermo_fac by value after reading the temperature back on the host; here the thermostat
 * state lives in device memory, d_nvt_state = float[4] {xi, eta, exp_thermo_fac, unused} (initial state {0, 0, 1, 0}), so
 * step one -> kinetic energy -> advanceThermostat -> force -> step two is one asynchronous chain.
 *   step one : v = (v + (1/2) a dt) * exp_fac ; pos += dt v ; wrap           (TwoStepNVTMTK.cc:120-160)
 *   advance  : T = 2 KE / ndof of the half-step velocities (double sum, fixed two-pass order over n_partial_blocks CTAs),
 *              xi' = xi + (dt/2)/tau^2 (T/T0 - 1), xi = xi' + the same, eta += xi' dt, exp_fac = exp(-xi dt/2)
 *   step two : a = F/m ; v = v * exp_fac + (1/2) dt a                        (TwoStepNVTMTK.cc:297-340)
 * ndof = IntegratorTwoStep::getNDOF(group) (D N - D for one method on all particles, IntegratorTwoStep.cc:270-283). */
B200_API int b200_nvt_mtk_step_one(float* d_pos, float* d_vel, const float* d_accel, int* d_image,
                                   const unsigned int* d_group_members, unsigned int group_size, const b200_box* box,
                                   const float* d_nvt_state, float deltaT, b200_stream_t stream);
B200_API int b200_nvt_mtk_step_two(float* d_vel, float* d_accel, const unsigned int* d_group_members,
                                   unsigned int group_size, const float* d_net_force, const float* d_nvt_state,
                                   float deltaT, b200_stream_t stream);
B200_API int b200_nvt_mtk_advance(float* d_nvt_state, double* d_partial, unsigned int n_partial_blocks, const float* d_vel,
                                  const unsigned int* d_group_members, unsigned int group_size, unsigned int ndof,
                                  float T, float tau, float deltaT, b200_stream_t stream);

/* -----------------------------------------------------------------------------------------------------------------------
 * TwoStepNPTMTK (Martyna-Tobias-Klein barostat + Nose-Hoover thermostat, point particles) -- replaces
 * gpu_npt_mtk_step_one / gpu_npt_mtk_wrap / gpu_npt_mtk_step_two / gpu_npt_mtk_rescale
 * (hoomd/md/TwoStepNPTMTKGPU.cuh:20-58, called from hoomd/md/TwoStepNPTMTKGPU.cc).  Same operands: the three upper-
 * triangular propagator matrices {xx, xy, xz, yy, yz, zz} are HOST arrays of 6 values (TwoStepNPTMTK::updatePropagator,
 * hoomd/md/TwoStepNPTMTK.cc:644-748, computed by the caller), exp_thermo_fac = exp(-(xi + mtk) dt / 2).  Step one moves the
 * group (velocity half kick, barostat matrix, thermostat factor, position scale + drift); the caller then wraps ALL particles
 * into the new box with b200_npt_mtk_wrap.  rescale_all != 0: the caller scales every position with b200_npt_mtk_rescale
 * before step one, which then skips the position scale (TwoStepNPTMTK.cc:251-270).  Per-particle arithmetic follows the
 * reference's CPU loops operation by operation. */
B200_API int b200_npt_mtk_step_one(float* d_pos, float* d_vel, const float* d_accel, const unsigned int* d_group_members,
                                   unsigned int group_size, float exp_thermo_fac, const float* mat_exp_v,
                                   const float* mat_exp_r, const float* mat_exp_r_int, float deltaT, int rescale_all,
                                   b200_stream_t stream);
B200_API int b200_npt_mtk_wrap(unsigned int N, float* d_pos, int* d_image, const b200_box* box, b200_stream_t stream);
B200_API int b200_npt_mtk_step_two(float* d_vel, float* d_accel, const unsigned int* d_group_members,
                                   unsigned int group_size, const float* d_net_force, const float* mat_exp_v, float deltaT,
                                   float exp_thermo_fac, b200_stream_t stream);
B200_API int b200_npt_mtk_rescale(unsigned int N, float* d_pos, float mat_exp_r_xx, float mat_exp_r_xy, float mat_exp_r_xz,
                                  float mat_exp_r_yy, float mat_exp_r_yz, float mat_exp_r_zz, b200_stream_t stream);

/* -----------------------------------------------------------------------------------------------------------------------
 * BoxResizeUpdater::update, particle part (hoomd/BoxResizeUpdater.cc:98-131; the reference does this on the host):
 * scale_particles != 0: pos = new_box.makeCoordinates(old_box.makeFraction(pos)); else wrap pos into new_box and update
 * the image flags.  The caller sets the new global box afterwards (ParticleData::setGlobalBox) -- the cell list then
 * re-dimensions itself (CellList.cc:151-171) and the neighbor list's distance check sees lambda = L_new / L_last_build
 * (b200_nlist_needs_update_check's lambda arguments, NeighborList.cc:981-1007). */
B200_API int b200_box_resize_particles(float* d_pos, int* d_image, unsigned int N, const b200_box* old_box,
                                       const b200_box* new_box, int scale_particles, b200_stream_t stream);

/* n draws of Saru(seed1,seed2,seed3).u32() and .s<float>(-1,1) per triple, evaluated on the device
 * (hoomd/Saru.h:324-340,417-441,603-640) -- used by the RNG parity tests. */
B200_API int b200_saru_draw(unsigned int* d_u32_out, float* d_f_out, const unsigned int* d_seeds3,
                            unsigned int ntriples, unsigned int ndraws, b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------------------------
 * SFCPackUpdater -- replaces gpu_generate_sorted_order / gpu_apply_sorted_order (hoomd/SFCPackUpdaterGPU.cuh).
 * Order = 3D Hilbert curve over a grid^3 binning of the fractional coordinates (hoomd/SFCPackUpdater.cc:112-250,
 * 490-583), ties broken by particle index (stable).
 * ------------------------------------------------------------------------------------------------------------- */
B200_API int b200_sfc_generate_sorted_order(unsigned int* d_sort_order, unsigned int* d_keys_scratch,
                                            unsigned int* d_scratch2x, void* d_temp, uint64_t temp_bytes,
                                            uint64_t* temp_bytes_needed, const float* d_pos, unsigned int N,
                                            const b200_box* box, unsigned int grid, b200_stream_t stream);

typedef struct b200_pdata_arrays /* the per-particle arrays of hoomd/ParticleData.h:1118-1165 that the path carries */
    {
    float* pos; float* vel; float* accel; int* image; unsigned int* tag; float* net_force; float* charge;
    float* diameter; unsigned int* body;     /* any pointer may be NULL = not carried */
    } b200_pdata_arrays;

/* out[i] = in[order[i]] for every carried array, rtag[tag[i]] = i  (hoomd/SFCPackUpdater.cc:111-243) */
B200_API int b200_sfc_apply_sorted_order(const b200_pdata_arrays* out, const b200_pdata_arrays* in,
                                         unsigned int* d_rtag, const unsigned int* d_sort_order, unsigned int N,
                                         b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------------------------
 * ComputeThermo -- replaces gpu_compute_thermo (hoomd/ComputeThermoGPU.cu:45-560): double-precision sums of
 * KE = sum m v^2/2, PE = sum net_force.w, W = sum (vxx+vyy+vzz) of the net virial.  d_out: double[3+6]
 * (ke, pe, virial trace, then the 6 pressure-tensor virial sums).  Two-pass, fixed order => deterministic.
 * ------------------------------------------------------------------------------------------------------------- */
B200_API int b200_compute_thermo(double* d_out, double* d_partial, unsigned int n_partial_blocks, const float* d_vel,
                                 const float* d_net_force, const float* d_net_virial, unsigned int virial_pitch,
                                 const unsigned int* d_group_members, unsigned int group_size, b200_stream_t stream);

/* ComputeThermo with pdata_flag::pressure_tensor set (hoomd/ComputeThermo.cc:224-240,300-318; gpu_compute_thermo with
 * compute_pressure_tensor, hoomd/ComputeThermoGPU.cu): d_out = double[13] {sum m vx vx, m vx vy, m vx vz, m vy vy, m vy vz,
 * m vz vz, sum of net_force.w, sum of the six virial rows xx, xy, xz, yy, yz, zz}; d_partial = double[13 * n_partial_blocks].
 * The caller forms KE = (xx + yy + zz) / 2 and P_ab = (kinetic_ab + virial_ab) / V as ComputeThermo.cc:352-372 does. */
B200_API int b200_compute_thermo_tensor(double* d_out, double* d_partial, unsigned int n_partial_blocks, const float* d_vel,
                                        const float* d_net_force, const float* d_net_virial, unsigned int virial_pitch,
                                        const unsigned int* d_group_members, unsigned int group_size, b200_stream_t stream);

/* b200_compute_thermo WITHOUT a caller scratch buffer and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same kernels, same fixed summation order => the same bits as
 * b200_compute_thermo with the same n_partial_blocks.  The partial sums live in a file-scope `static __device__` array of
 * B200_THERMO_STATIC_MAX_BLOCKS x 9 doubles (72 KiB, held by every GPU that loads the library until the process exits);
 * its device address is looked up with cudaGetSymbolAddress on the first call per device (with that device current --
 * make that first call outside a stream capture).  The call itself launches two kernels on `stream` and calls no
 * cudaMalloc / cudaFree / allocating library routine.  n_partial_blocks outside [1, B200_THERMO_STATIC_MAX_BLOCKS] is
 * rejected: cudaErrorInvalidValue (1) and a message on stderr.  EVERY call on a device shares the one array (it keeps the
 * last call's partial sums; every row that is read is overwritten first, so nothing needs clearing): calls that overlap
 * on the GPU -- different streams, different host threads -- race; calls ordered on one stream are safe. */
#define B200_THERMO_STATIC_MAX_BLOCKS 1024u
B200_API int b200_compute_thermo_static(double* d_out, unsigned int n_partial_blocks, const float* d_vel,
                                        const float* d_net_force, const float* d_net_virial, unsigned int virial_pitch,
                                        const unsigned int* d_group_members, unsigned int group_size,
                                        b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------------------------
 * ENGINE: device-resident system with the reference's object graph and step semantics
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct b200_system b200_system;

/* hoomd.context.initialize + SystemDefinition(N, box, ntypes): one system per process / GPU (device = cuda ordinal) */
B200_API b200_system* b200_sys_create(unsigned int N, const b200_box* box, unsigned int ntypes, int device);
B200_API void b200_sys_destroy(b200_system* s);
B200_API const char* b200_sys_last_error(b200_system* s);

/* Particle state in index order (tag == index at creation), HOST pointers; NULL = leave unchanged.
 * pos4 = (x,y,z,type bits), vel4 = (vx,vy,vz,mass), image3 int.  Forces a neighbor list rebuild. */
B200_API int b200_sys_set_particles(b200_system* s, const float* pos4, const float* vel4, const int* image3);
B200_API int b200_sys_set_accel(b200_system* s, const float* accel3);  /* + notifyAccelSet() */
/* Current state in INDEX order (use tag/rtag to key comparisons); any pointer may be NULL */
B200_API int b200_sys_get_particles(b200_system* s, float* pos4, float* vel4, float* accel3, int* image3,
                                    unsigned int* tag, unsigned int* rtag);

/* md.nlist.cell(r_buff, check_period) + md.pair.lj(r_cut, nlist) + pair_coeff + set_params(mode).
 * rcut/lj1/lj2/ron: ntypes x ntypes row-major (ron may be NULL). */
B200_API int b200_sys_setup_lj(b200_system* s, float r_buff, unsigned int check_period, const float* rcut,
                               const float* lj1, const float* lj2, const float* ron, int shift_mode);
/* md.pair.<evaluator> in place of md.pair.lj: params4 = ntypes x ntypes x 4 floats (the evaluator's param_type padded to
 * four components, meaning as listed at b200_compute_pair_forces); the cell-tile pair mode is LJ only */
B200_API int b200_sys_setup_pair(b200_system* s, int evaluator, float r_buff, unsigned int check_period,
                                 const float* rcut, const float* params4, const float* ron, int shift_mode);
/* BondData of the snapshot (hoomd/BondedGroupData.h): nbonds bonds between particle TAGS tags2[2b], tags2[2b+1] of bond
 * type types[b] < n_bond_types (HOST pointers).  Replaces any previous bond set. */
B200_API int b200_sys_set_bonds(b200_system* s, unsigned int nbonds, const unsigned int* tags2,
                                const unsigned int* types, unsigned int n_bond_types);
/* md.bond.fene() + bond_coeff.set per bond type (K, r0, lj1, lj2 arrays of n_bond_types; hoomd/md/bond.py:308-372), or
 * md.bond.harmonic() (K, r0; lj1 = lj2 = NULL, kind = 1).  kind 0 = FENE.  The bond force is a second ForceCompute of
 * the integrator, summed after the pair force (Integrator::computeNetForce, hoomd/Integrator.cc:308-420). */
B200_API int b200_sys_setup_bond_force(b200_system* s, int kind, const float* K, const float* r0, const float* lj1,
                                       const float* lj2);
/* nlist.reset_exclusions(exclusions=['bond']) (NeighborList::addExclusionsFromBonds, hoomd/md/NeighborList.cc:603-630);
 * on = 0: clearExclusions.  The md.nlist default is ['bond','body'] (hoomd/md/nlist.py:104-110). */
B200_API int b200_sys_nlist_exclusions_from_bonds(b200_system* s, int on);
/* per-ForceCompute arrays after b200_sys_compute_forces: which = 0 pair force, 1 bond force (index order; virial6 [6*N]) */
B200_API int b200_sys_get_force_component(b200_system* s, int which, float* force4, float* virial6);
/* pdata flags pressure_tensor | isotropic_virial (hoomd/md/PotentialPair.h:342-343) */
B200_API int b200_sys_set_virial(b200_system* s, int on);
/* md.integrate.mode_standard(dt) */
B200_API int b200_sys_setup_integrator(b200_system* s, float dt);
/* md.integrate.nve(group=all, limit, zero_force) */
B200_API int b200_sys_add_nve(b200_system* s, int limit, float limit_val, int zero_force);
/* md.integrate.langevin(group=all, kT, seed) + set_gamma per type (gamma NULL = 1.0) */
B200_API int b200_sys_add_langevin(b200_system* s, float kT, unsigned int seed, const float* gamma, int noiseless_t);
/* md.integrate.langevin(..., dscale=lambda, tally=...) (hoomd/md/integrate.py:844-940): use_lambda: gamma = lambda * diameter
 * instead of the per-type gamma (TwoStepLangevinBase.cc:29-70); tally: keep the reservoir energy (TwoStepLangevin::setTally).
 * Call after b200_sys_add_langevin; resets the tally to zero. */
B200_API int b200_sys_set_langevin_options(b200_system* s, int use_lambda, float lambda, int tally);
/* ParticleData diameters by particle TAG (HOST pointer, N floats; default 1.0) -- read by langevin(dscale=...) */
B200_API int b200_sys_set_diameters(b200_system* s, const float* diameter_by_tag);
/* the "langevin_reservoir_energy" log quantity (TwoStepLangevin::getLogValue, hoomd/md/TwoStepLangevin.cc:61-80,416-417) */
B200_API int b200_sys_get_langevin_reservoir_energy(b200_system* s, double* out);
/* hoomd.group.*: the ParticleGroup the integration method acts on, as n member TAGS (HOST pointer); n = 0: group.all().
 * The reference integrates m_group->getIndexArray() (hoomd/md/TwoStepNVEGPU.cc:48-70); non-members keep position and
 * velocity, forces on them are still computed; NDOF = 3 * members - 3 (IntegratorTwoStep::getNDOF). */
B200_API int b200_sys_set_group(b200_system* s, const unsigned int* member_tags, unsigned int n);
/* langevin / nvt set_params(kT=...), or a temperature variant evaluated by the caller per step (setT of the method) */
B200_API int b200_sys_set_temperature(b200_system* s, float kT);
/* md.nlist.cell(dist_check=False): rebuild when the check period is hit instead of when a particle moved r_buff/2
 * (NeighborList::setEvery(every, dist_check), hoomd/md/NeighborList.cc:1126-1135); default on */
B200_API int b200_sys_set_dist_check(b200_system* s, int on);
/* md.integrate.nvt(group=all, kT, tau): TwoStepNVTMTK with its own ComputeThermo; the thermostat state stays on the device
 * and advances inside the single-kernel MD step.  get_nvt_state: {xi, eta} (the method's IntegratorVariables). */
B200_API int b200_sys_add_nvt(b200_system* s, float kT, float tau);
B200_API int b200_sys_get_nvt_state(b200_system* s, float* xi_eta);
B200_API int b200_sys_set_nvt_state(b200_system* s, const float* xi_eta);
/* md.integrate.npt(group, kT, tau, P, tauP, couple, x, y, z, nph, rescale_all, gamma): TwoStepNPTMTK
 * (hoomd/md/TwoStepNPTMTK.cc) on an orthorhombic box.  couple: 0 none, 1 xy, 2 xz, 3 yz, 4 xyz; flags: baro_x 1 | baro_y 2 |
 * baro_z 4 (the tilt degrees of freedom need triclinic boxes and are rejected).  The barostat / thermostat variables and the
 * propagator stay on the host as in the reference; the steps are host-driven (b200_npt_mtk_* + b200_compute_thermo_tensor).
 * get_npt_state: out12 = IntegratorVariables {eta, xi, nuxx, nuxy, nuxz, nuyy, nuyz, nuzz, xi_rot, eta_rot} and the log
 * quantities {npt_thermostat_energy, npt_barostat_energy}; get_pressure_tensor: {xx, xy, xz, yy, yz, zz, P, KE}. */
/* md.pair.dpd(r_cut, nlist, kT, seed): turns the conservative DPD pair force set up with b200_sys_setup_pair(B200_PAIR_DPD, ...)
 * into the DPD THERMOSTAT (drag + random pair forces; integrate with md.integrate.nve).  The force depends on the
 * half-step velocities of both particles, so steps are host-driven (step one -> force -> step two as separate launches).
 * kT <= 0 switches back to the conservative force. */
B200_API int b200_sys_set_dpd_thermostat(b200_system* s, float kT, unsigned int seed);
B200_API int b200_sys_add_npt(b200_system* s, float kT, float tau, float P, float tauP, int couple, unsigned int flags,
                              int nph, int rescale_all, float gamma);
B200_API int b200_sys_get_npt_state(b200_system* s, float* out12);
B200_API int b200_sys_set_npt_state(b200_system* s, const float* vars10);
B200_API int b200_sys_get_pressure_tensor(b200_system* s, float* out8);
/* hoomd.update.box_resize at a period boundary: one BoxResizeUpdater::update towards the orthorhombic box (Lx, Ly, Lz).
 * Steps run host-driven until the next neighbor-list build (the reference's lambda-scaled distance check decides it), then
 * device-driven again in the new box. */
B200_API int b200_sys_set_box(b200_system* s, float Lx, float Ly, float Lz, int scale_particles);
B200_API void b200_sys_get_box(b200_system* s, float* L3);
/* hoomd.update.sort period (0 = disabled) */
B200_API int b200_sys_set_sorter(b200_system* s, unsigned int period);
B200_API int b200_sys_sort_now(b200_system* s);
B200_API void b200_sys_set_timestep(b200_system* s, unsigned int ts);
B200_API unsigned int b200_sys_get_timestep(b200_system* s);
/* Launch tuning: threads per particle of the pair kernel (0 = auto), fused device-driven stepping on/off */
B200_API int b200_sys_set_tuning(b200_system* s, unsigned int pair_tpp, int use_graph);
/* Pair-force path of the step loop.  mode 0: walk the reference-layout neighbor list (head list + global indices,
 * hoomd/md/PotentialPairGPU.cuh:100-330).  mode 1: cell tiles -- the particles are kept in CellList order (re-ordered at
 * every list rebuild), each CTA stages its cells + one halo layer in shared memory and the list stores 16-bit tile slots
 * (DESIGN.md section 4b); needs one GPU, a periodic orthorhombic box and >= 3 cells per axis, else an error.  mode -1
 * (default): the engine's choice -- always mode 0: on the current tree the tile path is slower at every size measured
 * (LJ 1M: 3.5k vs 4.4k timesteps/s, DESIGN.md section 4b has the captures), it is kept as an opt-in experiment.  bx,by,bz = cells per tile (0 = auto), tpp = lanes per particle (0 = auto, 1, 2, 4).
 * Same pairs, same rebuild decisions; forces differ by FP32 summation order only. */
B200_API int b200_sys_set_pair_mode(b200_system* s, int mode, unsigned int bx, unsigned int by, unsigned int bz,
                                    unsigned int tpp);
/* out8 = {tile path active, bx, by, bz, staged positions per CTA, tpp, row capacity, number of tiles} */
B200_API int b200_sys_get_pair_mode(b200_system* s, unsigned int* out8);

/* NeighborList::compute(ts) / ForceCompute::compute(ts) in isolation (kernel-level parity) */
B200_API int b200_sys_compute_nlist(b200_system* s, unsigned int ts);
B200_API int b200_sys_compute_forces(b200_system* s, unsigned int ts);
/* System::run(nsteps): prepRun, then per step updaters (sorter) + IntegratorTwoStep::update.  Blocks until done.
 * *elapsed_s = wall seconds of the step loop (excluding prepRun), like System.cc:622-637; *gpu_ms = CUDA-event time. */
B200_API int b200_sys_run(b200_system* s, unsigned int nsteps, double* elapsed_s, double* gpu_ms);
/* Isolated half steps on the current state (host net_force4 may be NULL = use the device net force) */
B200_API int b200_sys_step_one(b200_system* s, unsigned int ts);
B200_API int b200_sys_step_two(b200_system* s, unsigned int ts, const float* net_force4);

B200_API int b200_sys_get_forces(b200_system* s, float* force4, float* virial6 /* [6*N], pitch N */);
B200_API unsigned int b200_sys_get_nlist_size(b200_system* s);
B200_API int b200_sys_get_nlist(b200_system* s, unsigned int* n_neigh, unsigned int* head, unsigned int* nlist,
                                unsigned int nlist_len);
/* out: {num updates (distance-check triggered), forced updates, nlist overflows, kernels launched in the last run,
 *       dim.x, dim.y, dim.z, Nmax[0]} */
B200_API int b200_sys_get_stats(b200_system* s, unsigned long long* out8);
/* How the last run stepped.  out4 = {stepping mode (set_tuning's use_graph), MD steps per graph unit (0: data dependent),
 * 1 when the multi-step kernel k_md_steps (several MD steps per launch, until the next list rebuild; plain LJ systems whose
 * whole grid is co-resident) drives the graph, graph units launched by the last run}.
 * b200_sys_set_multistep(0) keeps the one-step kernel everywhere (A/B, debugging); nonzero = automatic (default). */
B200_API int b200_sys_get_stepping(b200_system* s, unsigned int* out4);
B200_API int b200_sys_set_multistep(b200_system* s, int mode);
/* Micro-benchmarks in the spirit of Compute::benchmark (hoomd/ForceCompute.cc:153-180, hoomd/md/NeighborList.cc:260-289):
 * launches kernel `which` (0 = pair force on the current list, 1 = cell list + neighbor list build, 3 / 4 = its cell-list /
 * neighbor-list half alone, 2 = the single-kernel MD step on a scratch copy of the integrator state, 5 = the same `iters` steps
 * inside ONE launch of the multi-step kernel: per-step time incl. its grid-wide barrier; error when it does not apply) `iters` times on the system's stream between two CUDA events and
 * returns the average milliseconds per launch.  State is not advanced. */
B200_API int b200_sys_time_kernel(b200_system* s, int which, unsigned int iters, double* avg_ms);
/* ComputeThermo at the current step: out = {KE, PE, temperature, pressure, ndof} */
B200_API int b200_sys_get_thermo(b200_system* s, double* out5);

/* ---------------------------------------------------------------------------------------------------------------
 * Multi-GPU: DomainDecomposition + Communicator(GPU) for one NVSwitch box, one process per GPU.
 * Replaces hoomd/DomainDecomposition.cc (findDecomposition :323-372 is host arithmetic done by the caller,
 * calculateLocalBox :531-557) and hoomd/CommunicatorGPU.cc (migrateParticles :1495, exchangeGhosts :1804,
 * beginUpdateGhosts/finishUpdateGhosts :2500/:2870).  Every rank exports ONE device slab (positions, tags, migration
 * mailboxes, flag words) as a CUDA IPC handle; the caller all-gathers the handles (torch.distributed / MPI / anything)
 * and hands them back; from then on ghosts, migrants and barrier flags travel as plain stores over NVLink from inside
 * the step kernels.  rank = (iz*grid[1] + iy)*grid[0] + ix.  One particle type in this round.
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct b200_mg_params
    {
    unsigned int world, rank;
    unsigned int grid[3];        /* ranks per axis (product == world) */
    b200_box     global_box;
    unsigned int N_global;       /* total particle count (capacity estimates) */
    unsigned int N_local;        /* particles this rank starts with */
    unsigned int Ncap;           /* local capacity, 0 = 1.25 * N_global/world + 8192 */
    float        r_ghost;        /* ghost layer width = max r_cut + r_buff (NeighborList::getGhostLayerWidth) */
    float        capacity_factor;/* ghost segment head room over the mean density, 0 = 1.5 */
    unsigned int ntypes;
    int          device;
    } b200_mg_params;

B200_API b200_system* b200_sys_create_mg(const b200_mg_params* p);
/* handle64: cudaIpcMemHandle_t of this rank's slab; local_box: DomainDecomposition::calculateLocalBox;
 * caps3 = {Ncap, Gcap, migrants per direction} */
B200_API int b200_sys_mg_export(b200_system* s, unsigned char* handle64, b200_box* local_box, unsigned int* caps3);
/* handles: world * 64 bytes in rank order (the own entry is ignored) */
B200_API int b200_sys_mg_connect(b200_system* s, const unsigned char* handles);
/* current local particle count (changes with migration); getters/setters move this many entries */
B200_API unsigned int b200_sys_get_n(b200_system* s);
/* global tags of the local particles (b200_sys_create numbers them 0..N-1) */
B200_API int b200_sys_set_tags(b200_system* s, const unsigned int* tags);
/* out: {N local, Ncap, Gcap, max N seen, max migrants/direction, max ghosts/direction, ghosts sent now, step barriers} */
B200_API int b200_sys_mg_stats(b200_system* s, unsigned long long* out8);
/* out10: {wait ns, count} of the four barrier kinds (0 = the per-step all-ranks barrier, 1..3 = the rebuild phases:
 * migration, ghost plan, overflow vote) measured on the device from arriving at the barrier to seeing every rank, then the
 * number of boundary chunks of k_md_step and the total number of chunks (Communicator profiling, hoomd/Communicator.cc) */
B200_API int b200_sys_mg_timing(b200_system* s, unsigned long long* out10);
/* this rank's ghost region (Gcap entries; unused slots hold x = 1e30): positions and global tags */
B200_API int b200_sys_mg_get_ghosts(b200_system* s, float* pos4, unsigned int* tag);

#ifdef __cplusplus
}
#endif
#endif /* HOOMD_B200_H_ */
