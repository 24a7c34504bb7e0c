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
    float *lj1, *lj2, *rcutsq, *ronsq;
    int shift_mode;
    /* Integrator */
    float dt; int method, limit, zero_force, noiseless_t; float limit_val, kT; unsigned int seed; float* gamma;
    unsigned int timestep;
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
