/* TEST INFRASTRUCTURE ONLY -- CPU restatement (plain C) of the reference's algorithm for the MD hot path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may load this.
 * Parity status: PINNED -- checked against the reference's own known-answer vectors (hoomd/md/test/test_lj_force.cc,
 * test_neighborlist.cc, test_nve_integrator.cc via tests/golden/) and against oracle/_ref/libhoomd_ref.so (the
 * unmodified reference CPU classes compiled from /root/reference) in tests/test_oracle.py.
 */
#ifndef ORACLE_H_
#define ORACLE_H_
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_system orc_system;

/* hoomd/Saru.h:324-340,417-441: n draws of u32() and of s<float>(-1,1) */
void orc_saru(unsigned int s1, unsigned int s2, unsigned int s3, unsigned int n, unsigned int* u32_out, float* f_out);

/* SystemDefinition(N, BoxDim(Lx,Ly,Lz), ntypes); tag == index */
orc_system* orc_create(unsigned int N, float Lx, float Ly, float Lz, unsigned int ntypes);
void orc_destroy(orc_system* s);
const char* orc_last_error(orc_system* s);
void orc_set_threads(orc_system* s, int nthreads); /* OpenMP threads for the nlist / force loops (1 = the reference's serial loop order) */
void orc_set_particles(orc_system* s, const float* pos4, const float* vel4, const int* image3);
void orc_set_accel(orc_system* s, const float* accel3);
void orc_get_particles(orc_system* s, float* pos4, float* vel4, float* accel3, int* image3, unsigned int* tag);
/* md.nlist.cell + md.pair.lj; matrices ntypes x ntypes row-major; full storage mode */
int orc_setup_lj(orc_system* s, float r_buff, unsigned int check_period, const float* rcut, const float* lj1,
                 const float* lj2, const float* ron, int shift_mode);
/* BondData + md.bond.fene (kind 0: K, r0, lj1, lj2) / md.bond.harmonic (kind 1: K, r0) + nlist bond exclusions:
 * hoomd/md/PotentialBond.h:155-300, EvaluatorBondFENE.h:95-133, EvaluatorBondHarmonic.h:75-92, NeighborList.cc:603-630,
 * 1275-1328 */
int orc_set_bonds(orc_system* s, unsigned int nbonds, const unsigned int* tags2, const unsigned int* types,
                  unsigned int n_bond_types);
int orc_setup_bond(orc_system* s, int kind, const float* K, const float* r0, const float* lj1, const float* lj2);
int orc_exclusions_from_bonds(orc_system* s, int on);
void orc_get_force_component(orc_system* s, int which, float* force4, float* virial6);
/* md.pair.<evaluator> (0 lj, 1 gauss, 2 yukawa, 3 mie, 4 force_shifted_lj, 5 morse), params4 = ntypes x ntypes x 4 floats */
int orc_setup_pair(orc_system* s, int evaluator, float r_buff, unsigned int check_period, const float* rcut,
                   const float* params4, const float* ron, int shift_mode);
void orc_set_virial(orc_system* s, int on);
int orc_setup_integrator(orc_system* s, float dt);
int orc_add_nve(orc_system* s, int limit, float limit_val, int zero_force);
int orc_add_langevin(orc_system* s, float kT, unsigned int seed, const float* gamma, int noiseless_t);
void orc_set_dist_check(orc_system* s, int on);
void orc_set_kT(orc_system* s, float kT);
int orc_add_nvt(orc_system* s, float kT, float tau);
int orc_set_box(orc_system* s, float Lx, float Ly, float Lz, int scale_particles);
void orc_get_box(orc_system* s, float* L3);
void orc_get_nvt_state(orc_system* s, float* xi_eta);
float orc_get_nvt_reservoir_energy(orc_system* s);
void orc_set_timestep(orc_system* s, unsigned int ts);
unsigned int orc_get_timestep(orc_system* s);
int orc_compute_nlist(orc_system* s, unsigned int ts);
int orc_compute_forces(orc_system* s, unsigned int ts);
int orc_run(orc_system* s, unsigned int nsteps, double* elapsed_s);
int orc_step_one(orc_system* s, unsigned int ts);
int orc_step_two(orc_system* s, unsigned int ts, const float* net_force4);
void orc_get_forces(orc_system* s, float* force4, float* virial6 /* [6*N] */);
/* parity-test scales: sums of absolute pair / bond contributions per particle, double[8*N] (see oracle.c) */
void orc_get_force_scales(orc_system* s, int which /* 1 pair, 2 bond, 3 both */, double* out8N);
unsigned int orc_get_nlist_size(orc_system* s);
void orc_get_nlist(orc_system* s, unsigned int* n_neigh, unsigned int* head, unsigned int* nlist, unsigned int len);
void orc_get_stats(orc_system* s, unsigned long long* out4); /* updates, forced, dim.x, Nmax[0] */
/* cell list in the reference's padded layout: out5 = dim x,y,z, Nmax, n_adj; arrays sized by the caller */
void orc_get_cell_dims(orc_system* s, unsigned int* out5);
void orc_get_cell_list(orc_system* s, unsigned int* cell_size, float* xyzf4);
/* ComputeThermo: {KE, PE, T, P, ndof} */
void orc_get_thermo(orc_system* s, double* out5);

#ifdef __cplusplus
}
#endif
#endif
