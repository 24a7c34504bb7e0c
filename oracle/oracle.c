/* TEST INFRASTRUCTURE ONLY -- see oracle.h.  Plain-C restatement of the reference's CPU algorithm for the MD hot path
 * (HOOMD-blue 2.3.0, SINGLE_PRECISION).  Compile with -ffp-contract=off (oracle/Makefile): every FP32 operation is
 * written in the reference's order so results are bit-identical to oracle/_ref for the serial (1-thread) case.
 * Each function cites the reference lines it follows (paths relative to /root/reference).
 */
#include "oracle.h"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { float x, y, z; } f3;
typedef struct { float x, y, z, w; } f4;
typedef struct { int x, y, z; } i3;

typedef struct /* BoxDim, hoomd/BoxDim.h:93-98,165-171 (orthorhombic) */
    {
    f3 lo, hi, L, Linv;
    unsigned char periodic[3];
    } box_t;

struct orc_system
    {
    unsigned int N, ntypes;
    box_t box;
    f4 *pos, *vel, *net_force, *last_pos;
    f3* accel;
    i3* image;
    unsigned int* tag;
    float* net_virial; /* 6 x N */
    int accel_set, compute_virial, nthreads;
    /* CellList */
    unsigned int dim[3], cl_Nmax, ncells;
    unsigned int* cell_size;
    f4* xyzf;
    unsigned int* cell_adj; unsigned int n_adj;
    int cl_first, cl_force; unsigned int cl_last;
    int no_dist_check;       /* NeighborList::m_dist_check == false (md.nlist.cell(dist_check=False)) */
    int cl_box_changed; f3 last_L; /* CellList::m_box_changed; NeighborList::m_last_L (nearest plane distances at the last build) */
    /* NeighborList */
    int have_nlist;
    float r_buff, rcut_max_max;
    float *r_cut, *r_listsq, *rcut_max;
    unsigned int *Nmax, *conditions, *head, *n_neigh, *nlist;
    size_t nlist_alloc;
    int nl_first, force_update, last_check_result;
    unsigned int nl_last, every, last_updated, last_checked;
    unsigned long long updates, forced;
    /* PotentialPairLJ */
    float *lj1, *lj2, *rcutsq, *ronsq; /* lj1/lj2 = the first two components of the evaluator's param_type */
    float *p3, *p4;                    /* third / fourth component (Mie, Morse) */
    int shift_mode, pair_kind;
    /* Integrator */
    float dt; int method, limit, zero_force, noiseless_t; float limit_val, kT; unsigned int seed; float* gamma;
    float nvt_tau, nvt_xi, nvt_eta, nvt_exp_fac; /* TwoStepNVTMTK: IntegratorVariables xi, eta and m_exp_thermo_fac */
    unsigned int timestep;
    /* BondData + PotentialBond + exclusions (tag == index in this restatement: no sorter) */
    unsigned int nbonds, n_bond_types;
    unsigned int *bond_a, *bond_b, *bond_type;
    int bond_kind; /* -1 none, 0 FENE, 1 harmonic */
    float *bond_K, *bond_r0, *bond_lj1, *bond_lj2;
    f4 *pair_force, *bond_force;
    float *pair_virial, *bond_virial;
    int exclusions_set; unsigned int ex_max; unsigned int *ex_n, *ex_list; /* ex_list[i*ex_max + k] */
    int bond_error;
    char error[256];
    };

/* ---- hoomd/Saru.h ---------------------------------------------------------------------------------------------- */
typedef struct { unsigned int x, y; } saru_t;

static saru_t saru_init(unsigned int seed1, unsigned int seed2, unsigned int seed3) /* Saru.h:324-340 */
    {
    saru_t s;
    seed3 ^= (seed1 << 7) ^ (seed2 >> 6);
    seed2 += (seed1 >> 4) ^ (seed3 >> 15);
    seed1 ^= (seed2 << 9) + (seed3 << 8);
    seed3 ^= 0xA5366B4Du * ((seed2 >> 11) ^ (seed1 << 1));
    seed2 += 0x72BE1579u * ((seed1 << 4) ^ (seed3 >> 16));
    seed1 ^= 0x3F38A6EDu * ((seed3 >> 5) ^ (unsigned int)(((int)seed2) >> 22));
    seed2 += seed1 * seed3;
    seed1 += seed3 ^ (seed2 >> 2);
    seed2 ^= (unsigned int)(((int)seed2) >> 17);
    s.x = 0x79dedea3u * (seed1 ^ (unsigned int)(((int)seed1) >> 14));
    s.y = (s.x + seed2) ^ (unsigned int)(((int)s.x) >> 8);
    s.x = s.x + (s.y * (s.y ^ 0xdddf97f5u));
    s.y = 0xABCB96F7u + (s.y >> 1);
    return s;
    }

static unsigned int saru_u32(saru_t* s) /* Saru.h:417-424,603-640 with steps = 1 */
    {
    s->x = 0x4beb5d59u * s->x + 0x2600e1f7u;
    s->y = s->y + 0x8009d14bu + ((unsigned int)(((int)s->y) >> 31) & 0xda879addu);
    {
    const unsigned int v = (s->x ^ (s->x >> 26)) + s->y;
    return (v ^ (v >> 20)) * 0x6957f5a7u;
    }
    }

static float saru_f(saru_t* s) /* Saru.h:436-441 */
    {
    const float TWO_N32 = 2.32830643653869628906250e-10f;
    return ((int)saru_u32(s)) * TWO_N32 + 0.5f;
    }

static float saru_s(saru_t* s, float a, float b) { return a + (b - a) * saru_f(s); } /* Saru.h:476-481,580-584 */

void orc_saru(unsigned int s1, unsigned int s2, unsigned int s3, unsigned int n, unsigned int* u32_out, float* f_out)
    {
    unsigned int i;
    if (u32_out)
        {
        saru_t s = saru_init(s1, s2, s3);
        for (i = 0; i < n; i++) u32_out[i] = saru_u32(&s);
        }
    if (f_out)
        {
        saru_t s = saru_init(s1, s2, s3);
        for (i = 0; i < n; i++) f_out[i] = saru_s(&s, -1.0f, 1.0f);
        }
    }

/* ---- hoomd/BoxDim.h ------------------------------------------------------------------------------------------------ */
static void box_init(box_t* b, float Lx, float Ly, float Lz)
    {
    b->L.x = Lx; b->L.y = Ly; b->L.z = Lz;
    b->hi.x = Lx / 2.0f; b->hi.y = Ly / 2.0f; b->hi.z = Lz / 2.0f;
    b->lo.x = -b->hi.x; b->lo.y = -b->hi.y; b->lo.z = -b->hi.z;
    b->Linv.x = 1.0f / Lx; b->Linv.y = 1.0f / Ly; b->Linv.z = 1.0f / Lz;
    b->periodic[0] = b->periodic[1] = b->periodic[2] = 1;
    }

static f3 box_make_fraction(const box_t* b, f3 v) /* BoxDim.h:238-245, ghost_width = 0, tilts = 0 */
    {
    f3 d, f;
    d.x = v.x - b->lo.x; d.y = v.y - b->lo.y; d.z = v.z - b->lo.z;
    f.x = (d.x * b->Linv.x + 0.0f) / (1.0f + 2.0f * 0.0f);
    f.y = (d.y * b->Linv.y + 0.0f) / (1.0f + 2.0f * 0.0f);
    f.z = (d.z * b->Linv.z + 0.0f) / (1.0f + 2.0f * 0.0f);
    return f;
    }

static f3 box_min_image(const box_t* b, f3 w) /* BoxDim.h:301-348 (host branch) */
    {
    if (b->periodic[2])
        {
        if (w.z >= b->hi.z) w.z -= b->L.z;
        else if (w.z < b->lo.z) w.z += b->L.z;
        }
    if (b->periodic[1])
        {
        if (w.y >= b->hi.y) { int i = (int)(w.y * b->Linv.y + 0.5f); w.y -= (float)i * b->L.y; }
        else if (w.y < b->lo.y) { int i = (int)(-w.y * b->Linv.y + 0.5f); w.y += (float)i * b->L.y; }
        }
    if (b->periodic[0])
        {
        if (w.x >= b->hi.x) { int i = (int)(w.x * b->Linv.x + 0.5f); w.x -= (float)i * b->L.x; }
        else if (w.x < b->lo.x) { int i = (int)(-w.x * b->Linv.x + 0.5f); w.x += (float)i * b->L.x; }
        }
    return w;
    }

static void box_wrap(const box_t* b, f4* w, i3* img) /* BoxDim.h:366-432, flags = 0 */
    {
    if (b->periodic[0])
        {
        if (w->x >= b->hi.x) { w->x -= b->L.x; img->x++; }
        else if (w->x < b->lo.x) { w->x += b->L.x; img->x--; }
        }
    if (b->periodic[1])
        {
        if (w->y >= b->hi.y) { w->y -= b->L.y; img->y++; }
        else if (w->y < b->lo.y) { w->y += b->L.y; img->y--; }
        }
    if (b->periodic[2])
        {
        if (w->z >= b->hi.z) { w->z -= b->L.z; img->z++; }
        else if (w->z < b->lo.z) { w->z += b->L.z; img->z--; }
        }
    }

static float dot3(f3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; } /* HOOMDMath.h:324-327 */
static unsigned int as_uint(float f) { unsigned int u; memcpy(&u, &f, 4); return u; }
static float as_float(unsigned int u) { float f; memcpy(&f, &u, 4); return f; }

/* ---- CellList (hoomd/CellList.cc) ----------------------------------------------------------------------------------- */
static void cl_dimensions(orc_system* s, float nominal_width) /* CellList.cc:69-131, no MPI */
    {
    s->dim[0] = (unsigned int)(s->box.L.x / nominal_width);
    s->dim[1] = (unsigned int)(s->box.L.y / nominal_width);
    s->dim[2] = (unsigned int)(s->box.L.z / nominal_width);
    for (int a = 0; a < 3; a++) if (s->dim[a] == 0) s->dim[a] = 1;
    s->ncells = s->dim[0] * s->dim[1] * s->dim[2];
    }

static int cmp_uint(const void* a, const void* b)
    {
    unsigned int x = *(const unsigned int*)a, y = *(const unsigned int*)b;
    return (x > y) - (x < y);
    }

static void cl_init_memory(orc_system* s) /* CellList.cc:276-436 */
    {
    unsigned int mx = s->dim[0], my = s->dim[1], mz = s->dim[2];
    if (s->cl_Nmax == 0)
        {
        s->cl_Nmax = (unsigned int)ceil((double)s->N / (double)s->ncells);
        if (s->cl_Nmax == 0) s->cl_Nmax = 1;
        }
    free(s->cell_size); free(s->xyzf); free(s->cell_adj);
    s->cell_size = (unsigned int*)calloc(s->ncells, sizeof(unsigned int));
    s->xyzf = (f4*)calloc((size_t)s->ncells * s->cl_Nmax, sizeof(f4));
    s->n_adj = (mx > 3 ? 3 : mx) * (my > 3 ? 3 : my) * (mz > 3 ? 3 : mz);
    s->cell_adj = (unsigned int*)calloc((size_t)s->ncells * s->n_adj, sizeof(unsigned int));
    for (int k = 0; k < (int)mz; k++)
        for (int j = 0; j < (int)my; j++)
            for (int i = 0; i < (int)mx; i++)
                {
                unsigned int cur = ((unsigned int)k * my + j) * mx + i, adj[27], n = 0, u = 0;
                for (int nk = k - 1; nk <= k + 1; nk++)
                    for (int nj = j - 1; nj <= j + 1; nj++)
                        for (int ni = i - 1; ni <= i + 1; ni++)
                            {
                            int wi = ni % (int)mx; if (wi < 0) wi += mx;
                            int wj = nj % (int)my; if (wj < 0) wj += my;
                            int wk = nk % (int)mz; if (wk < 0) wk += mz;
                            adj[n++] = ((unsigned int)wk * my + wj) * mx + wi;
                            }
                qsort(adj, n, sizeof(unsigned int), cmp_uint);
                for (unsigned int q = 0; q < n; q++)
                    if (q == 0 || adj[q] != adj[q - 1]) s->cell_adj[(size_t)cur * s->n_adj + u++] = adj[q];
                }
    }

/* CellList::computeCellList (CellList.cc:438-573); returns the overflow condition (conditions.x), -1/-2 on NaN / out of box */
static long cl_compute_once(orc_system* s)
    {
    unsigned int overflow = 0;
    memset(s->cell_size, 0, sizeof(unsigned int) * s->ncells);
    for (unsigned int n = 0; n < s->N; n++)
        {
        f3 p = {s->pos[n].x, s->pos[n].y, s->pos[n].z};
        if (isnan(p.x) || isnan(p.y) || isnan(p.z)) return -1;
        f3 f = box_make_fraction(&s->box, p);
        int ib = (int)(f.x * s->dim[0]), jb = (int)(f.y * s->dim[1]), kb = (int)(f.z * s->dim[2]);
        if ((f.x < -0.00001f || f.x >= 1.00001f) || (f.y < -0.00001f || f.y >= 1.00001f)
            || (f.z < -0.00001f || f.z >= 1.00001f))
            return -2;
        if (ib == (int)s->dim[0] && s->box.periodic[0]) ib = 0;
        if (jb == (int)s->dim[1] && s->box.periodic[1]) jb = 0;
        if (kb == (int)s->dim[2] && s->box.periodic[2]) kb = 0;
        if (ib < 0 || ib >= (int)s->dim[0] || jb < 0 || jb >= (int)s->dim[1] || kb < 0 || kb >= (int)s->dim[2])
            return -2;
        unsigned int bin = ((unsigned int)kb * s->dim[1] + jb) * s->dim[0] + ib;
        unsigned int offset = s->cell_size[bin];
        if (offset < s->cl_Nmax)
            {
            f4 e = {s->pos[n].x, s->pos[n].y, s->pos[n].z, as_float(n)};
            s->xyzf[(size_t)bin * s->cl_Nmax + offset] = e;
            }
        else if (offset + 1 > overflow)
            overflow = offset + 1;
        s->cell_size[bin]++;
        }
    return (long)overflow;
    }

static int cl_compute(orc_system* s, unsigned int ts) /* CellList::compute, CellList.cc:133-199 */
    {
    if (s->cl_box_changed) /* CellList.cc:151-171: same bin counts => only the width changes, else re-initialise */
        {
        const unsigned int od[3] = {s->dim[0], s->dim[1], s->dim[2]};
        cl_dimensions(s, s->rcut_max_max + s->r_buff);
        if ((od[0] != s->dim[0] || od[1] != s->dim[1] || od[2] != s->dim[2]) && s->cell_size) cl_init_memory(s);
        s->cl_box_changed = 0;
        s->cl_force = 1;
        }
    int should = s->cl_first || s->cl_force || s->cl_last != ts;
    if (!should) return 0;
    s->cl_first = 0; s->cl_force = 0; s->cl_last = ts;
    if (!s->cell_size) cl_init_memory(s);
    for (;;)
        {
        long c = cl_compute_once(s);
        if (c == -1) { snprintf(s->error, sizeof(s->error), "Error computing cell list: NaN position"); return 1; }
        if (c == -2) { snprintf(s->error, sizeof(s->error), "Error computing cell list: particle out of box"); return 1; }
        if ((unsigned int)c > s->cl_Nmax) { s->cl_Nmax = (unsigned int)c; cl_init_memory(s); }
        else break;
        }
    return 0;
    }

/* ---- NeighborList (hoomd/md/NeighborList.cc, NeighborListBinned.cc) ---------------------------------------------------- */
static void nl_update_rlist(orc_system* s) /* NeighborList.cc:354-400 */
    {
    unsigned int nt = s->ntypes;
    float r_cut_max = 0.0f;
    for (unsigned int i = 0; i < nt; i++)
        {
        float r_cut_max_i = 0.0f;
        for (unsigned int j = 0; j < nt; j++)
            {
            const float r_cut_ij = s->r_cut[i * nt + j];
            if (r_cut_ij > r_cut_max_i) r_cut_max_i = r_cut_ij;
            float r_list = (r_cut_ij > 0.0f) ? r_cut_ij + s->r_buff : 0.0f;
            s->r_listsq[i * nt + j] = r_list * r_list;
            }
        s->rcut_max[i] = r_cut_max_i;
        if (r_cut_max_i > r_cut_max) r_cut_max = r_cut_max_i;
        }
    s->rcut_max_max = r_cut_max;
    }

static void nl_build_head_list(orc_system* s) /* NeighborList.cc:1336-1380 */
    {
    unsigned int head = 0;
    for (unsigned int i = 0; i < s->N; i++)
        {
        s->head[i] = head;
        head += s->Nmax[as_uint(s->pos[i].w)];
        }
    if (head > s->nlist_alloc)
        {
        size_t alloc = s->nlist_alloc ? s->nlist_alloc : 1;
        while (head > alloc) alloc = ((size_t)(((float)alloc) * 1.125f)) + 1;
        s->nlist = (unsigned int*)realloc(s->nlist, alloc * sizeof(unsigned int));
        s->nlist_alloc = alloc;
        }
    }

static int nl_distance_check(orc_system* s) /* NeighborList.cc:964-1040; lambda = L_now / L_at_last_build per axis */
    {
    const f3 lambda = {s->box.L.x / s->last_L.x, s->box.L.y / s->last_L.y, s->box.L.z / s->last_L.z};
    float lambda_min = (lambda.x < lambda.y) ? lambda.x : lambda.y;
    lambda_min = (lambda_min < lambda.z) ? lambda_min : lambda.z;
    for (unsigned int i = 0; i < s->N; i++)
        {
        const float old_rmin = s->rcut_max[as_uint(s->pos[i].w)];
        const float rmax = old_rmin + s->r_buff;
        const float delta_max = (rmax * lambda_min - old_rmin) / 2.0f;
        const float maxsq = (delta_max > 0) ? delta_max * delta_max : 0;
        f3 dx = {s->pos[i].x - lambda.x * s->last_pos[i].x, s->pos[i].y - lambda.y * s->last_pos[i].y,
                 s->pos[i].z - lambda.z * s->last_pos[i].z};
        dx = box_min_image(&s->box, dx);
        if (dot3(dx) >= maxsq) return 1;
        }
    return 0;
    }

static int nl_needs_updating(orc_system* s, unsigned int ts) /* NeighborList.cc:1080-1163 */
    {
    int result = 0;
    if (s->last_checked == ts)
        {
        if (s->force_update) { s->force_update = 0; return 1; }
        return s->last_check_result;
        }
    s->last_checked = ts;
    if (!s->force_update && (ts < s->last_updated + s->every)) { s->last_check_result = 0; return 0; }
    if (s->force_update)
        {
        result = 1; s->force_update = 0; s->forced += 1; s->last_updated = ts;
        }
    else
        {
        /* distance checks off: rebuild when the check period is hit (or every step for period 0), NeighborList.cc:1126-1135 */
        if (s->r_buff < 1e-6f
            || (s->no_dist_check && (s->every == 0 || (s->every > 1 && ts == s->last_updated + s->every))))
            result = 1;
        else result = nl_distance_check(s);
        if (result) { s->last_updated = ts; s->updates += 1; }
        }
    s->last_check_result = result;
    return result;
    }

/* NeighborListBinned::buildNlist (NeighborListBinned.cc:79-240), full storage, no body filter, no diameter shift */
static int nl_build(orc_system* s, unsigned int ts)
    {
    if (cl_compute(s, ts)) return 1;
    {
    const float rmax = s->rcut_max_max + s->r_buff;
    if ((s->box.periodic[0] && s->box.L.x <= rmax * 2.0f) || (s->box.periodic[1] && s->box.L.y <= rmax * 2.0f)
        || (s->box.periodic[2] && s->box.L.z <= rmax * 2.0f))
        {
        snprintf(s->error, sizeof(s->error), "nlist: Simulation box is too small!");
        return 1;
        }
    }
    const unsigned int nt = s->ntypes;
#pragma omp parallel for schedule(static) num_threads(s->nthreads)
    for (int i = 0; i < (int)s->N; i++)
        {
        unsigned int cur_n_neigh = 0;
        const f3 my_pos = {s->pos[i].x, s->pos[i].y, s->pos[i].z};
        const unsigned int type_i = as_uint(s->pos[i].w);
        const unsigned int Nmax_i = s->Nmax[type_i];
        const unsigned int head_i = s->head[i];
        f3 f = box_make_fraction(&s->box, my_pos);
        int ib = (unsigned int)(f.x * s->dim[0]), jb = (unsigned int)(f.y * s->dim[1]), kb = (unsigned int)(f.z * s->dim[2]);
        if (ib == (int)s->dim[0] && s->box.periodic[0]) ib = 0;
        if (jb == (int)s->dim[1] && s->box.periodic[1]) jb = 0;
        if (kb == (int)s->dim[2] && s->box.periodic[2]) kb = 0;
        const unsigned int my_cell = ((unsigned int)kb * s->dim[1] + jb) * s->dim[0] + ib;
        for (unsigned int a = 0; a < s->n_adj; a++)
            {
            const unsigned int neigh_cell = s->cell_adj[(size_t)my_cell * s->n_adj + a];
            const unsigned int size = s->cell_size[neigh_cell];
            for (unsigned int o = 0; o < size; o++)
                {
                const f4 c = s->xyzf[(size_t)neigh_cell * s->cl_Nmax + o];
                const unsigned int j = as_uint(c.w);
                const unsigned int type_j = as_uint(s->pos[j].w);
                const float r_cut = s->r_cut[type_i * nt + type_j];
                if ((i == (int)j) || (r_cut <= 0.0f)) continue;
                f3 dx = {my_pos.x - c.x, my_pos.y - c.y, my_pos.z - c.z};
                dx = box_min_image(&s->box, dx);
                const float dr_sq = dot3(dx);
                if (dr_sq <= (s->r_listsq[type_i * nt + type_j] + 0.0f))
                    {
                    if (cur_n_neigh < Nmax_i) s->nlist[head_i + cur_n_neigh] = j;
                    else
                        {
#pragma omp critical
                        if (cur_n_neigh + 1 > s->conditions[type_i]) s->conditions[type_i] = cur_n_neigh + 1;
                        }
                    cur_n_neigh++;
                    }
                }
            }
        s->n_neigh[i] = cur_n_neigh;
        }
    return 0;
    }

static void nl_filter(orc_system* s) /* NeighborList::filterNlist, NeighborList.cc:1275-1328 */
    {
    for (unsigned int idx = 0; idx < s->N; idx++)
        {
        const unsigned int head = s->head[idx], n = s->n_neigh[idx], n_ex = s->ex_n[idx];
        unsigned int out = 0;
        for (unsigned int k = 0; k < n; k++)
            {
            const unsigned int cur = s->nlist[head + k];
            int excluded = 0;
            for (unsigned int e = 0; e < n_ex; e++)
                if (s->ex_list[(size_t)idx * s->ex_max + e] == cur) { excluded = 1; break; }
            if (!excluded) s->nlist[head + out++] = cur;
            }
        s->n_neigh[idx] = out;
        }
    }

static int nl_compute(orc_system* s, unsigned int ts) /* NeighborList::compute, NeighborList.cc:201-253 */
    {
    if (!s->have_nlist) { snprintf(s->error, sizeof(s->error), "nlist not set up"); return 1; }
    int should = s->nl_first || s->nl_last != ts;
    s->nl_first = 0; s->nl_last = ts;
    if (!should && !s->force_update) return 0;
    if (s->force_update) nl_build_head_list(s);
    if (nl_needs_updating(s, ts))
        {
        int overflowed;
        do
            {
            if (nl_build(s, ts)) return 1;
            overflowed = 0;
            for (unsigned int t = 0; t < s->ntypes; t++) /* checkConditions, NeighborList.cc:1389-1405 */
                if (s->conditions[t] > s->Nmax[t])
                    {
                    s->Nmax[t] = (s->conditions[t] > 8) ? (s->conditions[t] + 7) & ~7u : 8;
                    overflowed = 1;
                    }
            if (overflowed)
                {
                nl_build_head_list(s);
                memset(s->conditions, 0, sizeof(unsigned int) * s->ntypes);
                }
            } while (overflowed);
        if (s->exclusions_set) nl_filter(s);
        for (unsigned int i = 0; i < s->N; i++) /* setLastUpdatedPos, NeighborList.cc:1044-1066 */
            {
            f4 lp = {s->pos[i].x, s->pos[i].y, s->pos[i].z, 0.0f};
            s->last_pos[i] = lp;
            }
        s->last_L = s->box.L; /* NeighborList.cc:1062 (orthorhombic: nearest plane distance = L) */
        }
    return 0;
    }

/* one bond: dx = minImage(pos_b - pos_a), force_divr and the FULL bond energy; returns 1 on "bond out of bounds" */
static int bond_eval(orc_system* s, unsigned int b, f3* dx_out, float* force_divr_out, float* bond_eng_out)
    {
        const unsigned int ia = s->bond_a[b], ib = s->bond_b[b], t = s->bond_type[b]; /* rtag == identity */
        f3 dx = {s->pos[ib].x - s->pos[ia].x, s->pos[ib].y - s->pos[ia].y, s->pos[ib].z - s->pos[ia].z};
        dx = box_min_image(&s->box, dx);
        float rsq = dot3(dx);
        float force_divr = 0.0f, bond_eng = 0.0f;
        const float K = s->bond_K[t], r_0 = s->bond_r0[t];
        if (s->bond_kind == 0)
            {
            const float lj1 = s->bond_lj1[t], lj2 = s->bond_lj2[t];
            const float diameter_a = 1.0f, diameter_b = 1.0f; /* ParticleData default diameter */
            float rmdoverr = 1.0f;
            float rtemp = sqrtf(rsq) - diameter_a / 2 - diameter_b / 2 + 1.0f;
            rmdoverr = rtemp / sqrtf(rsq);
            rsq = rtemp * rtemp;
            {
            float r2inv = 1.0f / rsq;
            float r6inv = r2inv * r2inv * r2inv;
            float WCAforcemag_divr = 0.0f, pair_eng = 0.0f;
            float sigma6inv = lj2 / lj1;
            float epsilon = lj2 * lj2 / 4.0f / lj1;
            if (r6inv > sigma6inv / 2.0f)
                {
                WCAforcemag_divr = r2inv * r6inv * (12.0f * lj1 * r6inv - 6.0f * lj2);
                pair_eng = (r6inv * (lj1 * r6inv - lj2) + epsilon);
                }
            force_divr = WCAforcemag_divr * rmdoverr;
            bond_eng = pair_eng;
            if (K != 0.0f)
                {
                if (rsq >= r_0 * r_0)
                    {
                    snprintf(s->error, sizeof(s->error), "bond.fene: bond out of bounds");
                    return 1;
                    }
                force_divr += -K / (1.0f - rsq / (r_0 * r_0)) * rmdoverr;
                bond_eng += -0.5f * K * (r_0 * r_0) * logf(1.0f - rsq / (r_0 * r_0));
                }
            }
            }
        else
            {
            float r = sqrtf(rsq);
            force_divr = K * (r_0 / r - 1.0f);
            if (!isfinite(force_divr)) force_divr = 0.0f;
            bond_eng = 0.5f * K * (r_0 - r) * (r_0 - r);
            }
    *dx_out = dx;
    *force_divr_out = force_divr;
    *bond_eng_out = bond_eng;
    return 0;
    }

/* ---- PotentialBond<evaluator>::computeForces (PotentialBond.h:155-300) + EvaluatorBondFENE.h:95-133 /
 * EvaluatorBondHarmonic.h:75-92, then Integrator::computeNetForce (Integrator.cc:308-420): net = pair + bond ------------- */
static int compute_bond_forces(orc_system* s)
    {
    const unsigned int N = s->N;
    const int compute_virial = s->compute_virial;
    memcpy(s->pair_force, s->net_force, sizeof(f4) * N);
    memcpy(s->pair_virial, s->net_virial, sizeof(float) * 6 * (size_t)N);
    memset(s->bond_force, 0, sizeof(f4) * N);
    memset(s->bond_virial, 0, sizeof(float) * 6 * (size_t)N);
    for (unsigned int b = 0; b < s->nbonds; b++)
        {
        const unsigned int ia = s->bond_a[b], ib = s->bond_b[b];
        f3 dx;
        float force_divr = 0.0f, bond_eng = 0.0f;
        if (bond_eval(s, b, &dx, &force_divr, &bond_eng)) return 1;
        bond_eng *= 0.5f;
        {
        float bv[6] = {0, 0, 0, 0, 0, 0};
        if (compute_virial)
            {
            float force_div2r = (float)(1.0 / 2.0) * force_divr;
            bv[0] = dx.x * dx.x * force_div2r; bv[1] = dx.x * dx.y * force_div2r; bv[2] = dx.x * dx.z * force_div2r;
            bv[3] = dx.y * dx.y * force_div2r; bv[4] = dx.y * dx.z * force_div2r; bv[5] = dx.z * dx.z * force_div2r;
            }
        s->bond_force[ib].x += force_divr * dx.x; s->bond_force[ib].y += force_divr * dx.y;
        s->bond_force[ib].z += force_divr * dx.z; s->bond_force[ib].w += bond_eng;
        s->bond_force[ia].x -= force_divr * dx.x; s->bond_force[ia].y -= force_divr * dx.y;
        s->bond_force[ia].z -= force_divr * dx.z; s->bond_force[ia].w += bond_eng;
        if (compute_virial)
            for (int c = 0; c < 6; c++)
                {
                s->bond_virial[c * (size_t)N + ib] += bv[c];
                s->bond_virial[c * (size_t)N + ia] += bv[c];
                }
        }
        }
    for (unsigned int i = 0; i < N; i++)
        {
        s->net_force[i].x += s->bond_force[i].x; s->net_force[i].y += s->bond_force[i].y;
        s->net_force[i].z += s->bond_force[i].z; s->net_force[i].w += s->bond_force[i].w;
        }
    if (compute_virial)
        for (size_t k = 0; k < 6 * (size_t)N; k++) s->net_virial[k] += s->bond_virial[k];
    return 0;
    }

/* ---- EvaluatorPair*::evalForceAndEnergy: LJ (EvaluatorPairLJ.h:137-158), Gauss (EvaluatorPairGauss.h:100-122), Yukawa
 * (EvaluatorPairYukawa.h:98-126), Mie (EvaluatorPairMie.h:104-128), force-shifted LJ (EvaluatorPairForceShiftedLJ.h:97-127),
 * Morse (EvaluatorPairMorse.h:96-120); fast::exp / rsqrt / sqrt are libm calls in the CPU build (HOOMDMath.h) ------------- */
static int pair_eval(int kind, float rsq, float rcutsq, float p1, float p2, float p3, float p4, int energy_shift,
                     float* force_divr_out, float* pair_eng_out)
    {
    float force_divr = 0.0f, pair_eng = 0.0f;
    switch (kind)
        {
        case 0: /* LJ: (lj1, lj2) */
            {
            const float lj1 = p1, lj2 = p2;
            if (!(rsq < rcutsq && lj1 != 0)) return 0;
            float r2inv = 1.0f / rsq;
            float r6inv = r2inv * r2inv * r2inv;
            force_divr = r2inv * r6inv * (12.0f * lj1 * r6inv - 6.0f * lj2);
            pair_eng = r6inv * (lj1 * r6inv - lj2);
            if (energy_shift)
                {
                float rcut2inv = 1.0f / rcutsq;
                float rcut6inv = rcut2inv * rcut2inv * rcut2inv;
                pair_eng -= rcut6inv * (lj1 * rcut6inv - lj2);
                }
            break;
            }
        case 1: /* Gauss: (epsilon, sigma) */
            {
            const float epsilon = p1, sigma = p2;
            if (!(rsq < rcutsq)) return 0;
            float sigma_sq = sigma * sigma;
            float r_over_sigma_sq = rsq / sigma_sq;
            float exp_val = expf(-1.0f / 2.0f * r_over_sigma_sq);
            force_divr = epsilon / sigma_sq * exp_val;
            pair_eng = epsilon * exp_val;
            if (energy_shift) pair_eng -= epsilon * expf(-1.0f / 2.0f * rcutsq / sigma_sq);
            break;
            }
        case 2: /* Yukawa: (epsilon, kappa) */
            {
            const float epsilon = p1, kappa = p2;
            if (!(rsq < rcutsq && epsilon != 0)) return 0;
            float rinv = 1.0f / sqrtf(rsq);
            float r = 1.0f / rinv;
            float r2inv = 1.0f / rsq;
            float exp_val = expf(-kappa * r);
            force_divr = epsilon * exp_val * r2inv * (rinv + kappa);
            pair_eng = epsilon * exp_val * rinv;
            if (energy_shift)
                {
                float rcutinv = 1.0f / sqrtf(rcutsq);
                float rcut = 1.0f / rcutinv;
                pair_eng -= epsilon * expf(-kappa * rcut) * rcutinv;
                }
            break;
            }
        case 3: /* Mie: (mie1, mie2, n, m) */
            {
            const float mie1 = p1, mie2 = p2, mie3 = p3, mie4 = p4;
            if (!(rsq < rcutsq && mie1 != 0)) return 0;
            float r2inv = 1.0f / rsq;
            float rninv = powf(r2inv, mie3 / 2.0f);
            float rminv = powf(r2inv, mie4 / 2.0f);
            force_divr = r2inv * (mie3 * mie1 * rninv - mie4 * mie2 * rminv);
            pair_eng = mie1 * rninv - mie2 * rminv;
            if (energy_shift)
                {
                float rcutninv = 1.0f / powf(rcutsq, mie3 / 2.0f);
                float rcutminv = 1.0f / powf(rcutsq, mie4 / 2.0f);
                pair_eng -= mie1 * rcutninv - mie2 * rcutminv;
                }
            break;
            }
        case 4: /* force-shifted LJ: (lj1, lj2) */
            {
            const float lj1 = p1, lj2 = p2;
            if (!(rsq < rcutsq && lj1 != 0)) return 0;
            float r2inv = 1.0f / rsq;
            float r6inv = r2inv * r2inv * r2inv;
            force_divr = r2inv * r6inv * (12.0f * lj1 * r6inv - 6.0f * lj2);
            pair_eng = r6inv * (lj1 * r6inv - lj2);
            float rcut2inv = 1.0f / rcutsq;
            float rcut6inv = rcut2inv * rcut2inv * rcut2inv;
            if (energy_shift) pair_eng -= rcut6inv * (lj1 * rcut6inv - lj2);
            float rcut_r_inv = 1.0f / sqrtf(rsq * rcutsq);
            float force_rcut_at_rcut = rcut6inv * (12.0f * lj1 * rcut6inv - 6.0f * lj2);
            force_divr -= rcut_r_inv * force_rcut_at_rcut;
            pair_eng += (rsq * rcut_r_inv - 1.0f) * force_rcut_at_rcut;
            break;
            }
        case 5: /* Morse: (D0, alpha, r0) */
            {
            const float D0 = p1, alpha = p2, r0 = p3;
            if (!(rsq < rcutsq)) return 0;
            float r = sqrtf(rsq);
            float Exp_factor = expf(-alpha * (r - r0));
            pair_eng = D0 * Exp_factor * (Exp_factor - 2.0f);
            force_divr = 2.0f * D0 * alpha * Exp_factor * (Exp_factor - 1.0f) / r;
            if (energy_shift)
                {
                float rcut = sqrtf(rcutsq);
                float Exp_factor_cut = expf(-alpha * (rcut - r0));
                pair_eng -= D0 * Exp_factor_cut * (Exp_factor_cut - 2.0f);
                }
            break;
            }
        case 6: /* DPD conservative: (A, gamma), EvaluatorPairDPDThermo.h:139-160; no energy shift */
            {
            const float a = p1;
            if (!(rsq < rcutsq)) return 0;
            float rinv = 1.0f / sqrtf(rsq);
            float r = 1.0f / rinv;
            float rcutinv = 1.0f / sqrtf(rcutsq);
            float rcut = 1.0f / rcutinv;
            force_divr = a * (rinv - rcutinv);
            pair_eng = a * (rcut - r) - (1.0f / 2.0f) * a * rcutinv * (rcutsq - rsq);
            break;
            }
        default:
            return 0;
        }
    *force_divr_out = force_divr;
    *pair_eng_out = pair_eng;
    return 1;
    }

/* one listed pair (i, j) of PotentialPair::computeForces (PotentialPair.h:361-470): dx = minImage(pos_i - pos_j), the
 * evaluator, then the XPLOR smoothing; returns 0 when the pair does not contribute */
static int pair_ij(const orc_system* s, f3 pi, unsigned int typei, unsigned int j, f3* dx_out, float* force_divr_out,
                   float* pair_eng_out)
    {
    const unsigned int nt = s->ntypes;
    f3 dx = {pi.x - s->pos[j].x, pi.y - s->pos[j].y, pi.z - s->pos[j].z};
    const unsigned int typej = as_uint(s->pos[j].w);
    dx = box_min_image(&s->box, dx);
    const float rsq = dot3(dx);
    const unsigned int tp = typei * nt + typej;
    const float rcutsq = s->rcutsq[tp];
    float ronsq = 0.0f;
    int energy_shift = 0;
    if (s->shift_mode == 2) ronsq = s->ronsq[tp];
    if (s->shift_mode == 1) energy_shift = 1;
    else if (s->shift_mode == 2 && ronsq > rcutsq) energy_shift = 1;
    float force_divr = 0.0f, pair_eng = 0.0f;
    if (!pair_eval(s->pair_kind, rsq, rcutsq, s->lj1[tp], s->lj2[tp], s->p3 ? s->p3[tp] : 0.0f,
                   s->p4 ? s->p4[tp] : 0.0f, energy_shift, &force_divr, &pair_eng))
        return 0;
    if (s->shift_mode == 2 && rsq >= ronsq && rsq < rcutsq)
        {
        float old_pair_eng = pair_eng, old_force_divr = force_divr;
        float xplor_denom_inv = 1.0f / ((rcutsq - ronsq) * (rcutsq - ronsq) * (rcutsq - ronsq));
        float rsq_minus_r_cut_sq = rsq - rcutsq;
        float sw = rsq_minus_r_cut_sq * rsq_minus_r_cut_sq * (rcutsq + 2.0f * rsq - 3.0f * ronsq) * xplor_denom_inv;
        float ds_dr_divr = 12.0f * (rsq - ronsq) * rsq_minus_r_cut_sq * xplor_denom_inv;
        pair_eng = old_pair_eng * sw;
        force_divr = sw * old_force_divr - ds_dr_divr * old_pair_eng;
        }
    *dx_out = dx;
    *force_divr_out = force_divr;
    *pair_eng_out = pair_eng;
    return 1;
    }

/* ---- PotentialPair<EvaluatorPairLJ>::computeForces (PotentialPair.h:309-522, EvaluatorPairLJ.h:137-158), full list ------ */
static int compute_forces(orc_system* s, unsigned int ts)
    {
    if (nl_compute(s, ts)) return 1;
    const unsigned int N = s->N;
    const int compute_virial = s->compute_virial;
    memset(s->net_force, 0, sizeof(f4) * N);
    memset(s->net_virial, 0, sizeof(float) * 6 * (size_t)N);
#pragma omp parallel for schedule(static) num_threads(s->nthreads)
    for (int i = 0; i < (int)N; i++)
        {
        const f3 pi = {s->pos[i].x, s->pos[i].y, s->pos[i].z};
        const unsigned int typei = as_uint(s->pos[i].w);
        f3 fi = {0, 0, 0};
        float pei = 0.0f, vxx = 0, vxy = 0, vxz = 0, vyy = 0, vyz = 0, vzz = 0;
        const unsigned int head = s->head[i], size = s->n_neigh[i];
        for (unsigned int k = 0; k < size; k++)
            {
            f3 dx;
            float force_divr = 0.0f, pair_eng = 0.0f;
            if (pair_ij(s, pi, typei, s->nlist[head + k], &dx, &force_divr, &pair_eng))
                {
                {
                float force_div2r = force_divr * 0.5f;
                fi.x += dx.x * force_divr; fi.y += dx.y * force_divr; fi.z += dx.z * force_divr;
                pei += pair_eng * 0.5f;
                if (compute_virial)
                    {
                    vxx += force_div2r * dx.x * dx.x; vxy += force_div2r * dx.x * dx.y; vxz += force_div2r * dx.x * dx.z;
                    vyy += force_div2r * dx.y * dx.y; vyz += force_div2r * dx.y * dx.z; vzz += force_div2r * dx.z * dx.z;
                    }
                }
                }
            }
        s->net_force[i].x += fi.x; s->net_force[i].y += fi.y; s->net_force[i].z += fi.z; s->net_force[i].w += pei;
        if (compute_virial)
            {
            s->net_virial[0 * (size_t)N + i] += vxx; s->net_virial[1 * (size_t)N + i] += vxy;
            s->net_virial[2 * (size_t)N + i] += vxz; s->net_virial[3 * (size_t)N + i] += vyy;
            s->net_virial[4 * (size_t)N + i] += vyz; s->net_virial[5 * (size_t)N + i] += vzz;
            }
        }
    if (s->bond_kind >= 0) return compute_bond_forces(s);
    return 0;
    }

/* ---- TwoStepNVE / TwoStepLangevin (hoomd/md/TwoStepNVE.cc:77-122,255-300; TwoStepLangevin.cc:86-125,240-330) ------------ */
/* TwoStepNVTMTK::advanceThermostat (hoomd/md/TwoStepNVTMTK.cc:387-404), translational part: the temperature of the
   half-step velocities from ComputeThermo (double sum of m v^2, FP32 property, ComputeThermo.cc:241-252 / .h:109-116) */
static void nvt_advance(orc_system* s)
    {
    double ke = 0.0;
    if (s->compute_virial) /* pdata_flag::pressure_tensor: KE = half the trace of the kinetic pressure tensor (:222-238) */
        {
        double xx = 0.0, yy = 0.0, zz = 0.0;
        for (unsigned int j = 0; j < s->N; j++)
            {
            const double mass = s->vel[j].w;
            xx += mass * ((double)s->vel[j].x * (double)s->vel[j].x);
            yy += mass * ((double)s->vel[j].y * (double)s->vel[j].y);
            zz += mass * ((double)s->vel[j].z * (double)s->vel[j].z);
            }
        ke = 0.5f * (xx + yy + zz);
        }
    else
        {
        for (unsigned int j = 0; j < s->N; j++)
            ke += (double)s->vel[j].w * ((double)s->vel[j].x * (double)s->vel[j].x
                                         + (double)s->vel[j].y * (double)s->vel[j].y
                                         + (double)s->vel[j].z * (double)s->vel[j].z);
        ke *= 0.5f;
        }
    const unsigned int ndof = 3u * s->N - 3u; /* IntegratorTwoStep::getNDOF, IntegratorTwoStep.cc:270-283 */
    const float curr_T = 2.0f / (float)ndof * (float)ke;
    /* m_T->getValue() is a double (hoomd/Variant.h): the bracket and the sum are evaluated in double, then stored as Scalar */
    const float dt = s->dt, tau = s->nvt_tau;
    const double dev = (double)curr_T / (double)s->kT - (double)1.0f;
    const float c = (1.0f / 2.0f) * dt / tau / tau;
    const float xi_prime = (float)((double)s->nvt_xi + (double)c * dev);
    s->nvt_xi = (float)((double)xi_prime + (double)c * dev);
    s->nvt_eta += xi_prime * dt;
    s->nvt_exp_fac = expf(-(1.0f / 2.0f) * s->nvt_xi * dt);
    }

/* TwoStepNVTMTK::integrateStepOne, point particles (hoomd/md/TwoStepNVTMTK.cc:107-160) */
static void nvt_step_one(orc_system* s)
    {
    const float dt = s->dt, fac = s->nvt_exp_fac;
    for (unsigned int j = 0; j < s->N; j++)
        {
        float vx = s->vel[j].x + (1.0f / 2.0f) * s->accel[j].x * dt;
        float vy = s->vel[j].y + (1.0f / 2.0f) * s->accel[j].y * dt;
        float vz = s->vel[j].z + (1.0f / 2.0f) * s->accel[j].z * dt;
        vx *= fac; vy *= fac; vz *= fac;
        s->pos[j].x += dt * vx; s->pos[j].y += dt * vy; s->pos[j].z += dt * vz;
        s->vel[j].x = vx; s->vel[j].y = vy; s->vel[j].z = vz;
        }
    for (unsigned int j = 0; j < s->N; j++) box_wrap(&s->box, &s->pos[j], &s->image[j]);
    nvt_advance(s);
    }

/* TwoStepNVTMTK::integrateStepTwo (hoomd/md/TwoStepNVTMTK.cc:297-340) */
static void nvt_step_two(orc_system* s)
    {
    const float dt = s->dt, fac = s->nvt_exp_fac;
    for (unsigned int j = 0; j < s->N; j++)
        {
        const float minv = 1.0f / s->vel[j].w;
        s->accel[j].x = s->net_force[j].x * minv; s->accel[j].y = s->net_force[j].y * minv;
        s->accel[j].z = s->net_force[j].z * minv;
        float vx = s->vel[j].x * fac, vy = s->vel[j].y * fac, vz = s->vel[j].z * fac;
        vx += (1.0f / 2.0f) * dt * s->accel[j].x;
        vy += (1.0f / 2.0f) * dt * s->accel[j].y;
        vz += (1.0f / 2.0f) * dt * s->accel[j].z;
        s->vel[j].x = vx; s->vel[j].y = vy; s->vel[j].z = vz;
        }
    }

static void step_one(orc_system* s)
    {
    const float dt = s->dt;
    const int nve = s->method == 1;
    if (s->method == 3) { nvt_step_one(s); return; }
    for (unsigned int j = 0; j < s->N; j++)
        {
        if (nve && s->zero_force) s->accel[j].x = s->accel[j].y = s->accel[j].z = 0.0f;
        float dx = s->vel[j].x * dt + (1.0f / 2.0f) * s->accel[j].x * dt * dt;
        float dy = s->vel[j].y * dt + (1.0f / 2.0f) * s->accel[j].y * dt * dt;
        float dz = s->vel[j].z * dt + (1.0f / 2.0f) * s->accel[j].z * dt * dt;
        if (nve && s->limit)
            {
            float len = sqrtf(dx * dx + dy * dy + dz * dz);
            if (len > s->limit_val)
                {
                dx = dx / len * s->limit_val;
                dy = dy / len * s->limit_val;
                dz = dz / len * s->limit_val;
                }
            }
        s->pos[j].x += dx; s->pos[j].y += dy; s->pos[j].z += dz;
        s->vel[j].x += (1.0f / 2.0f) * s->accel[j].x * dt;
        s->vel[j].y += (1.0f / 2.0f) * s->accel[j].y * dt;
        s->vel[j].z += (1.0f / 2.0f) * s->accel[j].z * dt;
        box_wrap(&s->box, &s->pos[j], &s->image[j]);
        }
    }

static void step_two(orc_system* s, unsigned int timestep)
    {
    const float dt = s->dt;
    if (s->method == 3) { nvt_step_two(s); return; }
    for (unsigned int j = 0; j < s->N; j++)
        {
        if (s->method == 2)
            {
            saru_t saru = saru_init(s->tag[j], timestep, s->seed);
            float rx = saru_s(&saru, -1, 1), ry = saru_s(&saru, -1, 1), rz = saru_s(&saru, -1, 1);
            float gamma = s->gamma[as_uint(s->pos[j].w)];
            float coeff = sqrtf(6.0f * gamma * s->kT / dt);
            if (s->noiseless_t) coeff = 0.0f;
            float bd_fx = rx * coeff - gamma * s->vel[j].x;
            float bd_fy = ry * coeff - gamma * s->vel[j].y;
            float bd_fz = rz * coeff - gamma * s->vel[j].z;
            float minv = 1.0f / s->vel[j].w;
            s->accel[j].x = (s->net_force[j].x + bd_fx) * minv;
            s->accel[j].y = (s->net_force[j].y + bd_fy) * minv;
            s->accel[j].z = (s->net_force[j].z + bd_fz) * minv;
            s->vel[j].x += (1.0f / 2.0f) * s->accel[j].x * dt;
            s->vel[j].y += (1.0f / 2.0f) * s->accel[j].y * dt;
            s->vel[j].z += (1.0f / 2.0f) * s->accel[j].z * dt;
            }
        else
            {
            if (s->zero_force) s->accel[j].x = s->accel[j].y = s->accel[j].z = 0.0f;
            else
                {
                float minv = 1.0f / s->vel[j].w;
                s->accel[j].x = s->net_force[j].x * minv;
                s->accel[j].y = s->net_force[j].y * minv;
                s->accel[j].z = s->net_force[j].z * minv;
                }
            s->vel[j].x += (1.0f / 2.0f) * s->accel[j].x * dt;
            s->vel[j].y += (1.0f / 2.0f) * s->accel[j].y * dt;
            s->vel[j].z += (1.0f / 2.0f) * s->accel[j].z * dt;
            if (s->limit)
                {
                float vel = sqrtf(s->vel[j].x * s->vel[j].x + s->vel[j].y * s->vel[j].y + s->vel[j].z * s->vel[j].z);
                if ((vel * dt) > s->limit_val)
                    {
                    s->vel[j].x = s->vel[j].x / vel * s->limit_val / dt;
                    s->vel[j].y = s->vel[j].y / vel * s->limit_val / dt;
                    s->vel[j].z = s->vel[j].z / vel * s->limit_val / dt;
                    }
                }
            }
        }
    }

/* ---- public API ---------------------------------------------------------------------------------------------------------- */
orc_system* orc_create(unsigned int N, float Lx, float Ly, float Lz, unsigned int ntypes)
    {
    orc_system* s = (orc_system*)calloc(1, sizeof(orc_system));
    s->N = N; s->ntypes = ntypes ? ntypes : 1; s->nthreads = 1;
    box_init(&s->box, Lx, Ly, Lz);
    s->pos = (f4*)calloc(N, sizeof(f4)); s->vel = (f4*)calloc(N, sizeof(f4)); s->net_force = (f4*)calloc(N, sizeof(f4));
    s->last_pos = (f4*)calloc(N, sizeof(f4)); s->accel = (f3*)calloc(N, sizeof(f3)); s->image = (i3*)calloc(N, sizeof(i3));
    s->tag = (unsigned int*)calloc(N, sizeof(unsigned int)); s->net_virial = (float*)calloc(6 * (size_t)N, sizeof(float));
    for (unsigned int i = 0; i < N; i++) { s->tag[i] = i; s->vel[i].w = 1.0f; }
    s->cl_first = s->cl_force = 1; s->nl_first = 1; s->force_update = 1;
    s->bond_kind = -1;
    return s;
    }

/* BondData: bonds between particle tags (tag == index here) */
int orc_set_bonds(orc_system* s, unsigned int nbonds, const unsigned int* tags2, const unsigned int* types,
                  unsigned int n_bond_types)
    {
    free(s->bond_a); free(s->bond_b); free(s->bond_type);
    s->nbonds = nbonds; s->n_bond_types = n_bond_types;
    s->bond_a = (unsigned int*)calloc(nbonds ? nbonds : 1, sizeof(unsigned int));
    s->bond_b = (unsigned int*)calloc(nbonds ? nbonds : 1, sizeof(unsigned int));
    s->bond_type = (unsigned int*)calloc(nbonds ? nbonds : 1, sizeof(unsigned int));
    for (unsigned int b = 0; b < nbonds; b++)
        {
        if (tags2[2 * b] >= s->N || tags2[2 * b + 1] >= s->N || types[b] >= n_bond_types)
            {
            snprintf(s->error, sizeof(s->error), "bond: tag or type out of bounds");
            return 1;
            }
        s->bond_a[b] = tags2[2 * b]; s->bond_b[b] = tags2[2 * b + 1]; s->bond_type[b] = types[b];
        }
    return 0;
    }

/* md.bond.fene (kind 0) / md.bond.harmonic (kind 1) */
int orc_setup_bond(orc_system* s, int kind, const float* K, const float* r0, const float* lj1, const float* lj2)
    {
    const unsigned int nb = s->n_bond_types, N = s->N;
    if (!nb) { snprintf(s->error, sizeof(s->error), "bond: No bonds are defined."); return 1; }
    free(s->bond_K); free(s->bond_r0); free(s->bond_lj1); free(s->bond_lj2);
    s->bond_K = (float*)calloc(nb, sizeof(float)); s->bond_r0 = (float*)calloc(nb, sizeof(float));
    s->bond_lj1 = (float*)calloc(nb, sizeof(float)); s->bond_lj2 = (float*)calloc(nb, sizeof(float));
    for (unsigned int t = 0; t < nb; t++)
        {
        s->bond_K[t] = K[t]; s->bond_r0[t] = r0[t];
        if (kind == 0) { s->bond_lj1[t] = lj1[t]; s->bond_lj2[t] = lj2[t]; }
        }
    if (!s->bond_force)
        {
        s->pair_force = (f4*)calloc(N, sizeof(f4)); s->bond_force = (f4*)calloc(N, sizeof(f4));
        s->pair_virial = (float*)calloc(6 * (size_t)N, sizeof(float));
        s->bond_virial = (float*)calloc(6 * (size_t)N, sizeof(float));
        }
    s->bond_kind = kind;
    return 0;
    }

/* NeighborList::addExclusionsFromBonds (NeighborList.cc:603-630) / clearExclusions; duplicates skipped (:431-470) */
int orc_exclusions_from_bonds(orc_system* s, int on)
    {
    const unsigned int N = s->N;
    free(s->ex_n); free(s->ex_list);
    s->ex_n = NULL; s->ex_list = NULL; s->ex_max = 0; s->exclusions_set = 0;
    s->force_update = 1;
    if (!on) return 0;
    s->ex_n = (unsigned int*)calloc(N, sizeof(unsigned int));
    unsigned int* cnt = (unsigned int*)calloc(N, sizeof(unsigned int));
    for (unsigned int b = 0; b < s->nbonds; b++) { cnt[s->bond_a[b]]++; cnt[s->bond_b[b]]++; }
    unsigned int mx = 1;
    for (unsigned int i = 0; i < N; i++) if (cnt[i] > mx) mx = cnt[i];
    free(cnt);
    s->ex_max = mx;
    s->ex_list = (unsigned int*)calloc((size_t)N * mx, sizeof(unsigned int));
    for (unsigned int b = 0; b < s->nbonds; b++)
        {
        const unsigned int ta = s->bond_a[b], tb = s->bond_b[b];
        int dup = 0;
        for (unsigned int e = 0; e < s->ex_n[ta]; e++) if (s->ex_list[(size_t)ta * mx + e] == tb) dup = 1;
        if (dup) continue;
        s->ex_list[(size_t)ta * mx + s->ex_n[ta]++] = tb;
        s->ex_list[(size_t)tb * mx + s->ex_n[tb]++] = ta;
        }
    s->exclusions_set = 1;
    return 0;
    }

/* which = 0 pair force, 1 bond force (after orc_compute_forces) */
void orc_get_force_component(orc_system* s, int which, float* force4, float* virial6)
    {
    const f4* f = s->bond_kind < 0 ? s->net_force : (which ? s->bond_force : s->pair_force);
    const float* v = s->bond_kind < 0 ? s->net_virial : (which ? s->bond_virial : s->pair_virial);
    if (force4) memcpy(force4, f, sizeof(f4) * s->N);
    if (virial6) memcpy(virial6, v, sizeof(float) * 6 * (size_t)s->N);
    }

void orc_destroy(orc_system* s)
    {
    if (!s) return;
    free(s->pos); free(s->vel); free(s->net_force); free(s->last_pos); free(s->accel); free(s->image); free(s->tag);
    free(s->net_virial); free(s->cell_size); free(s->xyzf); free(s->cell_adj); free(s->r_cut); free(s->r_listsq);
    free(s->rcut_max); free(s->Nmax); free(s->conditions); free(s->head); free(s->n_neigh); free(s->nlist);
    free(s->lj1); free(s->lj2); free(s->rcutsq); free(s->ronsq); free(s->gamma);
    free(s->bond_a); free(s->bond_b); free(s->bond_type); free(s->bond_K); free(s->bond_r0); free(s->bond_lj1);
    free(s->bond_lj2); free(s->pair_force); free(s->bond_force); free(s->pair_virial); free(s->bond_virial);
    free(s->ex_n); free(s->ex_list); free(s->p3); free(s->p4);
    free(s);
    }

const char* orc_last_error(orc_system* s) { return s->error; }
void orc_set_threads(orc_system* s, int n) { s->nthreads = n > 0 ? n : 1; }

void orc_set_particles(orc_system* s, const float* pos4, const float* vel4, const int* image3)
    {
    if (pos4) memcpy(s->pos, pos4, sizeof(f4) * s->N);
    if (vel4) memcpy(s->vel, vel4, sizeof(f4) * s->N);
    if (image3) memcpy(s->image, image3, sizeof(i3) * s->N);
    s->force_update = 1; s->cl_force = 1;
    }

void orc_set_accel(orc_system* s, const float* accel3) { memcpy(s->accel, accel3, sizeof(f3) * s->N); s->accel_set = 1; }

void orc_get_particles(orc_system* s, float* pos4, float* vel4, float* accel3, int* image3, unsigned int* tag)
    {
    if (pos4) memcpy(pos4, s->pos, sizeof(f4) * s->N);
    if (vel4) memcpy(vel4, s->vel, sizeof(f4) * s->N);
    if (accel3) memcpy(accel3, s->accel, sizeof(f3) * s->N);
    if (image3) memcpy(image3, s->image, sizeof(i3) * s->N);
    if (tag) memcpy(tag, s->tag, sizeof(unsigned int) * s->N);
    }

int orc_setup_lj(orc_system* s, float r_buff, unsigned int check_period, const float* rcut, const float* lj1,
                 const float* lj2, const float* ron, int shift_mode)
    {
    const unsigned int nt = s->ntypes, n2 = nt * nt;
    s->r_buff = r_buff; s->every = check_period; s->shift_mode = shift_mode;
    s->r_cut = (float*)calloc(n2, 4); s->r_listsq = (float*)calloc(n2, 4); s->rcut_max = (float*)calloc(nt, 4);
    s->lj1 = (float*)calloc(n2, 4); s->lj2 = (float*)calloc(n2, 4); s->rcutsq = (float*)calloc(n2, 4);
    s->ronsq = (float*)calloc(n2, 4);
    for (unsigned int k = 0; k < n2; k++)
        {
        s->r_cut[k] = rcut[k]; s->lj1[k] = lj1[k]; s->lj2[k] = lj2[k];
        s->rcutsq[k] = rcut[k] * rcut[k];                 /* PotentialPair::setRcut, PotentialPair.h:237-250 */
        s->ronsq[k] = ron ? ron[k] * ron[k] : 0.0f;
        }
    nl_update_rlist(s);
    s->Nmax = (unsigned int*)calloc(nt, 4); s->conditions = (unsigned int*)calloc(nt, 4);
    for (unsigned int t = 0; t < nt; t++) s->Nmax[t] = 8;  /* NeighborList.cc:92-98 */
    s->head = (unsigned int*)calloc(s->N, 4); s->n_neigh = (unsigned int*)calloc(s->N, 4);
    s->nlist_alloc = 8 * (size_t)s->N; s->nlist = (unsigned int*)calloc(s->nlist_alloc ? s->nlist_alloc : 1, 4);
    cl_dimensions(s, s->rcut_max_max + s->r_buff);          /* NeighborListBinned.cc:56-76 */
    s->last_L = s->box.L;                                   /* NeighborList.cc:58 */
    s->have_nlist = 1; s->force_update = 1; s->cl_force = 1; s->nl_first = 1;
    s->pair_kind = 0;
    return 0;
    }

/* md.pair.<evaluator>: 0 lj, 1 gauss, 2 yukawa, 3 mie, 4 force_shifted_lj, 5 morse, 6 dpd_conservative; params4 = ntypes x ntypes x 4 */
int orc_setup_pair(orc_system* s, int evaluator, float r_buff, unsigned int check_period, const float* rcut,
                   const float* params4, const float* ron, int shift_mode)
    {
    const unsigned int n2 = s->ntypes * s->ntypes;
    float* a = (float*)calloc(n2, 4);
    float* b = (float*)calloc(n2, 4);
    free(s->p3); free(s->p4);
    s->p3 = (float*)calloc(n2, 4); s->p4 = (float*)calloc(n2, 4);
    for (unsigned int k = 0; k < n2; k++)
        {
        a[k] = params4[4 * k]; b[k] = params4[4 * k + 1]; s->p3[k] = params4[4 * k + 2]; s->p4[k] = params4[4 * k + 3];
        }
    int rc = orc_setup_lj(s, r_buff, check_period, rcut, a, b, ron, shift_mode);
    free(a); free(b);
    if (evaluator < 0 || evaluator > 6) { snprintf(s->error, sizeof(s->error), "unknown pair evaluator"); return 1; }
    s->pair_kind = evaluator;
    return rc;
    }

void orc_set_virial(orc_system* s, int on) { s->compute_virial = on; }
int orc_setup_integrator(orc_system* s, float dt) { s->dt = dt; return 0; }
int orc_add_nve(orc_system* s, int limit, float limit_val, int zero_force)
    {
    s->method = 1; s->limit = limit; s->limit_val = limit_val; s->zero_force = zero_force;
    return 0;
    }
int orc_add_langevin(orc_system* s, float kT, unsigned int seed, const float* gamma, int noiseless_t)
    {
    s->method = 2; s->limit = 0; s->zero_force = 0; s->kT = kT; s->noiseless_t = noiseless_t;
    seed = seed * 0x12345677u + 0x12345u; seed ^= (seed >> 16); seed *= 0x45679u; /* TwoStepLangevinBase.cc:54 */
    s->seed = seed;
    free(s->gamma);
    s->gamma = (float*)calloc(s->ntypes, 4);
    for (unsigned int t = 0; t < s->ntypes; t++) s->gamma[t] = gamma ? gamma[t] : 1.0f;
    return 0;
    }
int orc_add_nvt(orc_system* s, float kT, float tau) /* TwoStepNVTMTK ctor + initializeIntegratorVariables: xi = eta = 0 */
    {
    s->method = 3; s->limit = 0; s->zero_force = 0; s->kT = kT; s->nvt_tau = tau;
    s->nvt_xi = s->nvt_eta = 0.0f; s->nvt_exp_fac = 1.0f;
    return 0;
    }
void orc_set_dist_check(orc_system* s, int on) /* NeighborList::setEvery(every, dist_check): ends in forceUpdate() */
    {
    s->no_dist_check = !on;
    s->force_update = 1;
    }
void orc_set_kT(orc_system* s, float kT) { s->kT = kT; } /* TwoStepLangevinBase::setT / TwoStepNVTMTK::setT */
void orc_get_nvt_state(orc_system* s, float* xi_eta) { xi_eta[0] = s->nvt_xi; xi_eta[1] = s->nvt_eta; }
float orc_get_nvt_reservoir_energy(orc_system* s) /* TwoStepNVTMTK::getLogValue, TwoStepNVTMTK.cc:82-107 */
    {
    const float g = (float)(3u * s->N - 3u);
    const float xi = s->nvt_xi, eta = s->nvt_eta, tau = s->nvt_tau;
    return (float)((double)g * (double)s->kT * (double)(xi * xi * tau * tau / 2.0f + eta));
    }
/* BoxResizeUpdater::update for one target box (hoomd/BoxResizeUpdater.cc:60-136), orthorhombic */
int orc_set_box(orc_system* s, float Lx, float Ly, float Lz, int scale_particles)
    {
    const box_t cur = s->box;
    /* the tilt terms of no_change are 0/0 = NaN in the reference => `no_change` is false for every orthorhombic box */
    box_t nb = cur;
    nb.L.x = Lx; nb.L.y = Ly; nb.L.z = Lz;               /* BoxDim::setL, BoxDim.h:165-171 */
    nb.hi.x = Lx / 2.0f; nb.hi.y = Ly / 2.0f; nb.hi.z = Lz / 2.0f;
    nb.lo.x = -nb.hi.x; nb.lo.y = -nb.hi.y; nb.lo.z = -nb.hi.z;
    nb.Linv.x = 1.0f / Lx; nb.Linv.y = 1.0f / Ly; nb.Linv.z = 1.0f / Lz;
    s->box = nb;
    s->cl_box_changed = 1;
    if (scale_particles)
        for (unsigned int i = 0; i < s->N; i++)
            {
            f3 p = {s->pos[i].x, s->pos[i].y, s->pos[i].z};
            f3 f = box_make_fraction(&cur, p);
            s->pos[i].x = nb.lo.x + f.x * nb.L.x;          /* BoxDim::makeCoordinates, BoxDim.h:257-263 */
            s->pos[i].y = nb.lo.y + f.y * nb.L.y;
            s->pos[i].z = nb.lo.z + f.z * nb.L.z;
            }
    else
        for (unsigned int i = 0; i < s->N; i++) box_wrap(&s->box, &s->pos[i], &s->image[i]);
    return 0;
    }
void orc_get_box(orc_system* s, float* L3) { L3[0] = s->box.L.x; L3[1] = s->box.L.y; L3[2] = s->box.L.z; }
void orc_set_timestep(orc_system* s, unsigned int ts) { s->timestep = ts; }
unsigned int orc_get_timestep(orc_system* s) { return s->timestep; }
int orc_compute_nlist(orc_system* s, unsigned int ts) { return nl_compute(s, ts); }
int orc_compute_forces(orc_system* s, unsigned int ts) { return compute_forces(s, ts); }

static double now_s(void)
    {
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return t.tv_sec + 1e-9 * t.tv_nsec;
    }

int orc_run(orc_system* s, unsigned int nsteps, double* elapsed_s) /* System.cc:484-599, IntegratorTwoStep.cc:96-161,336-398 */
    {
    if (!s->method) { snprintf(s->error, sizeof(s->error), "no integration method"); return 1; }
    if (compute_forces(s, s->timestep)) return 1;
    if (!s->accel_set)
        {
        for (unsigned int j = 0; j < s->N; j++) /* Integrator::computeAccelerations, Integrator.cc:222-256 */
            {
            float minv = 1.0f / s->vel[j].w;
            s->accel[j].x = s->net_force[j].x * minv; s->accel[j].y = s->net_force[j].y * minv;
            s->accel[j].z = s->net_force[j].z * minv;
            }
        s->accel_set = 1;
        }
    double t0 = now_s();
    for (unsigned int k = 0; k < nsteps; k++)
        {
        step_one(s);
        if (compute_forces(s, s->timestep + 1)) return 1;
        step_two(s, s->timestep);
        s->timestep++;
        }
    if (elapsed_s) *elapsed_s = now_s() - t0;
    return 0;
    }

int orc_step_one(orc_system* s, unsigned int ts) { (void)ts; step_one(s); return 0; }
int orc_step_two(orc_system* s, unsigned int ts, const float* net_force4)
    {
    if (net_force4) memcpy(s->net_force, net_force4, sizeof(f4) * s->N);
    step_two(s, ts);
    return 0;
    }

void orc_get_forces(orc_system* s, float* force4, float* virial6)
    {
    if (force4) memcpy(force4, s->net_force, sizeof(f4) * s->N);
    if (virial6) memcpy(virial6, s->net_virial, sizeof(float) * 6 * (size_t)s->N);
    }

unsigned int orc_get_nlist_size(orc_system* s) { return (unsigned int)s->nlist_alloc; }
void orc_get_nlist(orc_system* s, unsigned int* n_neigh, unsigned int* head, unsigned int* nlist, unsigned int len)
    {
    if (n_neigh) memcpy(n_neigh, s->n_neigh, 4 * (size_t)s->N);
    if (head) memcpy(head, s->head, 4 * (size_t)s->N);
    if (nlist) memcpy(nlist, s->nlist, 4 * (len < s->nlist_alloc ? len : s->nlist_alloc));
    }
void orc_get_stats(orc_system* s, unsigned long long* out4)
    {
    out4[0] = s->updates; out4[1] = s->forced; out4[2] = s->dim[0]; out4[3] = s->Nmax ? s->Nmax[0] : 0;
    }
void orc_get_cell_dims(orc_system* s, unsigned int* out5)
    {
    out5[0] = s->dim[0]; out5[1] = s->dim[1]; out5[2] = s->dim[2]; out5[3] = s->cl_Nmax; out5[4] = s->n_adj;
    }
void orc_get_cell_list(orc_system* s, unsigned int* cell_size, float* xyzf4)
    {
    if (cell_size) memcpy(cell_size, s->cell_size, 4 * (size_t)s->ncells);
    if (xyzf4) memcpy(xyzf4, s->xyzf, sizeof(f4) * (size_t)s->ncells * s->cl_Nmax);
    }

/* Parity-test scale (not a reference quantity): per particle the sums of ABSOLUTE pair / bond contributions that make up
 * its force, energy and virial on the current neighbor list and bond set -- out[0*N+i] = sum_j |f_ij| (vector norm),
 * out[1*N+i] = sum_j |V_ij / 2|, out[(2+c)*N+i] = sum_j |virial_c term|, in double.  |a - b| <= 1e-5 * scale is the FP32
 * "1e-5 relative" bar of BASELINE.json for sums whose terms cancel.  Call after orc_compute_forces. */
void orc_get_force_scales(orc_system* s, int which /* 1 pair, 2 bond, 3 both */, double* out8N)
    {
    const unsigned int N = s->N;
    memset(out8N, 0, sizeof(double) * 8 * (size_t)N);
    for (unsigned int i = 0; (which & 1) && i < N; i++)
        {
        const f3 pi = {s->pos[i].x, s->pos[i].y, s->pos[i].z};
        const unsigned int typei = as_uint(s->pos[i].w);
        const unsigned int head = s->head[i], size = s->n_neigh[i];
        for (unsigned int k = 0; k < size; k++)
            {
            f3 dx;
            float force_divr = 0.0f, pair_eng = 0.0f;
            if (!pair_ij(s, pi, typei, s->nlist[head + k], &dx, &force_divr, &pair_eng)) continue;
            const double fd = fabs((double)force_divr);
            out8N[0 * (size_t)N + i] += fd * sqrt((double)dx.x * dx.x + (double)dx.y * dx.y + (double)dx.z * dx.z);
            out8N[1 * (size_t)N + i] += 0.5 * fabs((double)pair_eng);
            const double c[6] = {(double)dx.x * dx.x, (double)dx.x * dx.y, (double)dx.x * dx.z, (double)dx.y * dx.y,
                                 (double)dx.y * dx.z, (double)dx.z * dx.z};
            for (int q = 0; q < 6; q++) out8N[(2 + q) * (size_t)N + i] += 0.5 * fd * fabs(c[q]);
            }
        }
    if (s->bond_kind >= 0 && (which & 2))
        for (unsigned int b = 0; b < s->nbonds; b++)
            {
            f3 dx;
            float force_divr = 0.0f, bond_eng = 0.0f;
            if (bond_eval(s, b, &dx, &force_divr, &bond_eng)) continue;
            const double fd = fabs((double)force_divr);
            const double c[6] = {(double)dx.x * dx.x, (double)dx.x * dx.y, (double)dx.x * dx.z, (double)dx.y * dx.y,
                                 (double)dx.y * dx.z, (double)dx.z * dx.z};
            const unsigned int ab[2] = {s->bond_a[b], s->bond_b[b]};
            for (int e = 0; e < 2; e++)
                {
                const unsigned int i = ab[e];
                out8N[0 * (size_t)N + i] += fd * sqrt(c[0] + c[3] + c[5]);
                out8N[1 * (size_t)N + i] += 0.5 * fabs((double)bond_eng);
                for (int q = 0; q < 6; q++) out8N[(2 + q) * (size_t)N + i] += 0.5 * fd * fabs(c[q]);
                }
            }
    }

void orc_get_thermo(orc_system* s, double* out5) /* ComputeThermo.cc:195-380, FP32 properties */
    {
    double ke = 0.0, pe = 0.0, W = 0.0;
    for (unsigned int j = 0; j < s->N; j++)
        {
        ke += (double)s->vel[j].w * ((double)s->vel[j].x * (double)s->vel[j].x + (double)s->vel[j].y * (double)s->vel[j].y
                                     + (double)s->vel[j].z * (double)s->vel[j].z);
        pe += (double)s->net_force[j].w;
        }
    ke *= 0.5f;
    if (s->compute_virial)
        {
        double vxx = 0, vyy = 0, vzz = 0;
        for (unsigned int j = 0; j < s->N; j++)
            {
            vxx += (double)s->net_virial[0 * (size_t)s->N + j];
            vyy += (double)s->net_virial[3 * (size_t)s->N + j];
            vzz += (double)s->net_virial[5 * (size_t)s->N + j];
            }
        W = (float)(1. / 3.) * (vxx + vyy + vzz);
        }
    {
    const double ndof = 3.0 * s->N - 3.0;
    const float volume = s->box.L.x * s->box.L.y * s->box.L.z;
    out5[0] = (float)ke;
    out5[1] = (float)pe;
    out5[2] = (float)(2.0f / (float)ndof) * (float)ke;
    out5[3] = s->compute_virial ? (float)((2.0 * ke / 3.0f + W) / volume) : 0.0;
    out5[4] = ndof;
    }
    }
