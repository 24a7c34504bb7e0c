// Several MD steps per launch for SMALL systems (the whole grid co-resident): k_md_steps runs the single-kernel MD step in a
// loop until the neighbor list needs a rebuild or the chunk ends.  At 64k particles a step kernel runs 17 us and the gap
// between two graph nodes (conditional IF evaluation + kernel launch) 4-7 us: the loop replaces that gap by a grid-wide
// barrier built on the ticket the step already takes for its rebuild decision.
//
//   per step:   phase 1 (neighbor walk on buffer X) -> phase 2 (step two, step one into buffer Y, distance check)
//               -> ticket (a 128-byte line of its own); the LAST CTA takes NeighborList::needsUpdating for the next step
//               (decide_body on a register copy of the control block) and publishes ONE 8-byte release word that carries
//               the new epoch AND the next step's control block, replicated over 16 lines; every CTA spins on its copy,
//               takes one acquire fence, swaps X and Y.  Row metadata and the CTA's particle state stay on chip
//               (PersistRegs / PersistShared) from step to step.
//   exit:       rebuild wanted / chunk finished / abort.  The caller's rebuild bracket and the next launch always start from
//               buffer A (a.pos_in): a launch that stops with the current positions in B copies its own particles back.
//
// Arithmetic per particle is that of k_md_step (same bodies, same order): trajectories are bit-identical to the one-step
// kernel's.  Plain variants only: one particle type, no virial, LJ, NVE or Langevin without options, no second force, no
// group, single GPU -- everything else keeps the one-step kernel.  The launch is cooperative (co-residency guaranteed by
// the runtime); the epoch wait is bounded all the same, so a lost CTA turns into an error instead of a hung GPU.
#pragma once
#include "md_step_kernel.cuh"

namespace b200
{
constexpr unsigned int PERSIST_STOP_DONE = 1u, PERSIST_STOP_REBUILD = 2u, PERSIST_STOP_ERROR = 4u;

// -DB200_PERSIST_TRACE (tools/trace_persist.py, never in the shipped library): thread 0 of every CTA stamps clock64 at the
// phase boundaries of its first PERSIST_TRACE_STEPS steps into a device array that b200_debug_persist_trace copies out
#ifdef B200_PERSIST_TRACE
constexpr unsigned int PERSIST_TRACE_STEPS = 64, PERSIST_TRACE_MARKS = 8, PERSIST_TRACE_CTAS = 1536;
__device__ unsigned long long g_persist_trace[PERSIST_TRACE_CTAS * PERSIST_TRACE_STEPS * PERSIST_TRACE_MARKS];
__device__ unsigned int g_persist_trace_step[PERSIST_TRACE_CTAS];
#define PERSIST_MARK(m)                                                                                                     \
    if (threadIdx.x == 0 && blockIdx.x < PERSIST_TRACE_CTAS && g_persist_trace_step[blockIdx.x] < PERSIST_TRACE_STEPS)        \
        g_persist_trace[(blockIdx.x * PERSIST_TRACE_STEPS + g_persist_trace_step[blockIdx.x]) * PERSIST_TRACE_MARKS + (m)]   \
            = (unsigned long long)clock64();
#define PERSIST_MARK_END()                                                                                                  \
    if (threadIdx.x == 0 && blockIdx.x < PERSIST_TRACE_CTAS && g_persist_trace_step[blockIdx.x] < PERSIST_TRACE_STEPS)        \
        {                                                                                                                   \
        unsigned int smid;                                                                                                  \
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));                                                                   \
        g_persist_trace[(blockIdx.x * PERSIST_TRACE_STEPS + g_persist_trace_step[blockIdx.x]) * PERSIST_TRACE_MARKS + 7]     \
            = smid;                                                                                                         \
        g_persist_trace_step[blockIdx.x] += 1u;                                                                             \
        }
#else
#define PERSIST_MARK(m)
#define PERSIST_MARK_END()
#endif

// -DB200_FAULT_INJECT (tests/test_fault_gpu.py, never in the shipped library): CTA (g_fault_lose_cta - 1) skips its ticket
// once, so that the bounded wait on the release word -- and the host's handling of the error it raises -- can be exercised
#ifdef B200_FAULT_INJECT
__device__ unsigned int g_fault_lose_cta;
#endif

__device__ __forceinline__ unsigned long long* persist_ticket(StepState* st)
    {
    return reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(st) + STEP_SYNC_OFFSET);
    }
__device__ __forceinline__ unsigned long long* persist_release_word(StepState* st, unsigned int copy)
    {
    return reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(st) + STEP_SYNC_OFFSET
                                                 + (1u + copy % STEP_SYNC_COPIES) * STEP_SYNC_LINE);
    }
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p)
    {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
    }

// control words are read at GPU scope (a plain volatile load is system scope)
__device__ __forceinline__ unsigned int ld_volatile_u32(const unsigned int* p)
    {
    unsigned int v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
    }
// release / acquire fence at GPU scope (the acquire side also drops this SM's L1 lines: the next step's neighbor gathers
// must see the positions the other CTAs just stored)
__device__ __forceinline__ void persist_fence() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

// The control block of a step as the k_md_steps CTAs need it -- s_ctl[0] steps left (0, 1 = last, 2 = more), [1] ts_step_two,
// [2] store_force | box_scaled << 1, [3] stop reasons.  At launch it is read word by word; between the steps of a launch the
// LAST CTA of a step packs it, together with the new epoch, into ONE 8-byte release word and every CTA gets it with the very
// load it spins on: six dependent GPU-scope loads behind the barrier were 2 us of a 17 us step (profiles/r2y_trace_*).
// The release word is replicated over STEP_SYNC_COPIES 128-byte lines behind the step state (CTA b spins on copy b mod 16)
// and the ticket has a line of its own: with ticket, control block and release word in one line, the late CTAs' ticket
// atomics queued behind a thousand spinning loads at one L2 slice.
constexpr unsigned int CTL_STORE_FORCE = 1u, CTL_BOX_SCALED = 2u, CTL_REBUILD = 4u, CTL_ERROR = 8u, CTL_DONE = 16u,
                       CTL_LAST = 32u;

__device__ __forceinline__ void persist_read_ctl(StepState* st, unsigned int* s_ctl)
    {
    const unsigned int left = ld_volatile_u32(&st->steps_left);
    s_ctl[0] = left > 2u ? 2u : left;
    s_ctl[1] = ld_volatile_u32(&st->ts_step_two);
    s_ctl[2] = (ld_volatile_u32(&st->store_force) ? 1u : 0u) | (ld_volatile_u32(&st->box_scaled) ? 2u : 0u);
    s_ctl[3] = (ld_volatile_u32(&st->need_rebuild) ? PERSIST_STOP_REBUILD : 0u)
               | ((ld_volatile_u32(&st->abort) || ld_volatile_u32(&st->nvt_error)) ? PERSIST_STOP_ERROR : 0u);
    }

// last CTA of a step, after the rebuild decision: the next step's control block + the new epoch in one word
__device__ __forceinline__ unsigned long long persist_pack_ctl(const StepState* st, unsigned int epoch)
    {
    const unsigned int left = st->steps_left;
    const unsigned int flags = (st->store_force ? CTL_STORE_FORCE : 0u) | (st->box_scaled ? CTL_BOX_SCALED : 0u)
                               | (st->need_rebuild ? CTL_REBUILD : 0u) | ((st->abort || st->nvt_error) ? CTL_ERROR : 0u)
                               | (left == 0u ? CTL_DONE : 0u) | (left == 1u ? CTL_LAST : 0u);
    return ((unsigned long long)(((epoch & 0xffffffu) << 8) | flags) << 32) | st->ts_step_two;
    }

__device__ __forceinline__ void persist_unpack_ctl(unsigned long long w, unsigned int* s_ctl)
    {
    const unsigned int flags = (unsigned int)(w >> 32) & 0xffu;
    s_ctl[0] = (flags & CTL_DONE) ? 0u : ((flags & CTL_LAST) ? 1u : 2u);
    s_ctl[1] = (unsigned int)w;
    s_ctl[2] = flags & (CTL_STORE_FORCE | CTL_BOX_SCALED);
    s_ctl[3] = ((flags & CTL_REBUILD) ? PERSIST_STOP_REBUILD : 0u) | ((flags & CTL_ERROR) ? PERSIST_STOP_ERROR : 0u);
    }

// What a CTA keeps on chip (registers for the walker role, shared memory for the integrator role) from one step of a launch
// to the next.  Nothing else writes these while the launch runs:
// the neighbor list (row length, row start, this lane's first chunk of indices) and last_pos change only in the rebuild
// bracket between launches, and position / velocity / image of particle i are written by the thread that integrates i.
// With them a step starts its position gathers at once and its integration needs no global load at all: the dependent
// loads they replace were ~1.5 us of every CTA's 8 us of work per step (after each barrier the SM's L1 is empty, so every
// one of them was an L2 round trip).  Arithmetic and its order are unchanged: bit-identical trajectories.
struct PersistRegs // walker role: lane `sub` of particle slot threadIdx.x / TPP
    {
    unsigned int size, head;
    uint4 first;
    };
template<int PPB> struct PersistShared // integrator role: particle threadIdx.x of the CTA (shared memory: 68 B per particle)
    {
    float4 pos[PPB]; // also what the walkers of the next step read for their own particle
    float4 vel[PPB];
    float4 lp[PPB];  // last_pos.xyz, w = this particle's (type's) squared distance-check threshold
    int4 img[PPB];   // image.xyz, w = tag (Langevin)
    float gamma[PPB];
    float lam[3];            // lambda of the distance check (1 unless the box changed since the build) ...
    unsigned int row_scaled; // ... taken from the scaled (1) or the plain (0) row of the table
    };

template<class Evaluator, int PPB>
__device__ __forceinline__ void persist_load_check_row(const MdStepArgs<Evaluator>& a, PersistShared<PPB>& S, bool box_scaled,
                                                       bool mine)
    {
    const float* tab = a.maxsq + (box_scaled ? a.ntypes + 3u : 0u);
    if (mine) S.lp[threadIdx.x].w = tab[(unsigned int)__float_as_int(S.pos[threadIdx.x].w)];
    if (threadIdx.x < 3) S.lam[threadIdx.x] = tab[a.ntypes + threadIdx.x];
    if (threadIdx.x == 3) S.row_scaled = box_scaled ? 1u : 0u;
    }

template<class Evaluator, int TPP>
__device__ __forceinline__ void persist_init(const MdStepArgs<Evaluator>& a, PersistRegs& R,
                                             PersistShared<MD_BLOCK / TPP>& S, unsigned int ppb, bool box_scaled)
    {
    const unsigned int N = a.N;
        {
        const unsigned int slot = threadIdx.x / TPP, sub = threadIdx.x % TPP;
        const unsigned int i = blockIdx.x * ppb + slot;
        const bool active = slot < ppb && i < N;
        R.size = active ? a.n_neigh[i] : 0u;
        R.head = !active ? 0u : (a.nmax_uniform ? i * a.nmax_uniform : a.head_list[i]);
        R.first = make_uint4(0u, 0u, 0u, 0u);
        const bool aligned = (R.head & 3u) == 0u && (reinterpret_cast<unsigned long long>(a.nlist) & 15ull) == 0ull;
        if (active && aligned && sub < ((R.size + 3u) >> 2))
            R.first = ld_index4(reinterpret_cast<const uint4*>(a.nlist + R.head) + sub);
        }
    const unsigned int i = blockIdx.x * ppb + threadIdx.x;
    const bool mine = threadIdx.x < ppb && i < N;
    if (mine)
        {
        const float4 p = a.pos_in[i];
        S.pos[threadIdx.x] = p;
        S.vel[threadIdx.x] = a.vel[i];
        const int3 im = a.image[i];
        const float4 lp = a.last_pos[i];
        S.lp[threadIdx.x] = make_float4(lp.x, lp.y, lp.z, 0.0f);
        unsigned int tag = 0u;
        float gamma = 0.0f;
        if (a.method != 1)
            {
            tag = a.tag[i];
            gamma = a.gamma[(unsigned int)__float_as_int(p.w)];
            }
        S.img[threadIdx.x] = make_int4(im.x, im.y, im.z, (int)tag);
        S.gamma[threadIdx.x] = gamma;
        }
    persist_load_check_row(a, S, box_scaled, mine);
    __syncthreads();
    }

// one MD step reading buffer X = (ODD ? a.pos_out : a.pos_in), writing the other.  Returns 0 to go on, else the stop reason.
template<class Evaluator, int TPP, int SHIFT_MODE, bool ODD>
__device__ __forceinline__ unsigned int md_persist_step(const MdStepArgs<Evaluator>& a, StepState* st,
                                                        const PairTables<Evaluator>& T, float4* s_force,
                                                        PersistShared<MD_BLOCK / TPP>& S, unsigned int* s_ctl,
                                                        unsigned int& epoch, PersistRegs& R)
    {
    constexpr int PPB = MD_BLOCK / TPP;
    const float4* pin = ODD ? a.pos_out : a.pos_in;
    float4* pout = ODD ? const_cast<float4*>(a.pos_in) : a.pos_out;
    const unsigned int steps_left = s_ctl[0], ts2 = s_ctl[1];
    const bool box_scaled = (s_ctl[2] & 2u) != 0u;
    const bool last = steps_left == 1u;
    const bool store_force = last || (s_ctl[2] & 1u);
    const unsigned int N = a.N;
    // particles per CTA: PPB unless the launcher chose a larger, SM-count-aligned grid (HOOMD_B200_PERSIST_EVEN)
    const unsigned int ppb = (N + gridDim.x - 1u) / gridDim.x;

    if ((S.row_scaled != 0u) != box_scaled) // (never inside a launch today: the flag only falls with a rebuild decision)
        {
        __syncthreads();
        persist_load_check_row(a, S, box_scaled, threadIdx.x < ppb && blockIdx.x * ppb + threadIdx.x < N);
        __syncthreads();
        }
    PERSIST_MARK(0)
    // ---- phase 1 (as in k_md_step) ----
        {
        const unsigned int slot = threadIdx.x / TPP;
        const unsigned int i = blockIdx.x * ppb + slot;
        const unsigned int sub = threadIdx.x % TPP;
        const bool active = slot < ppb && i < N;
        const float4 pi4 = active ? S.pos[slot] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        const PairAccum r = pair_accumulate<Evaluator, TPP, SHIFT_MODE, false, false>(
            pi4, R.size, R.head, sub, pin, a.box, a.nlist, T, a.ntypes,
            !active || box_interior(pi4, a.box, a.wrap_skip_margin), &R.first);
        if (sub == 0) s_force[slot] = make_float4(r.fx, r.fy, r.fz, r.pe);
        if (threadIdx.x == 0) { PERSIST_MARK(1) }
        }
    __syncthreads();
    PERSIST_MARK(2)

    // ---- phase 2: md_step_integrate for the plain case ----
    bool moved = false;
    const unsigned int i = blockIdx.x * ppb + threadIdx.x;
    if (threadIdx.x < ppb && i < N)
        {
        const float4 F = s_force[threadIdx.x];
        const float4 pi4 = S.pos[threadIdx.x];
        float4 v = S.vel[threadIdx.x];
        if (store_force) a.net_force[i] = F;
        const bool nve = a.method == 1;
        float3s acc;
        if (nve)
            nve_step_two_body(v, acc, F, a.dt, a.limit != 0, a.limit_val, a.zero_force != 0);
        else
            langevin_step_two_body(v, acc, F, (unsigned int)S.img[threadIdx.x].w, ts2, a.seed, S.gamma[threadIdx.x], a.kT, a.dt,
                                   a.noiseless_t != 0, 3u);
        a.accel[i] = acc;
        if (!last)
            {
            float4 p = pi4;
            const int4 im4 = S.img[threadIdx.x];
            int3 img = make_int3(im4.x, im4.y, im4.z);
            nve_step_one_body(p, v, acc, img, a.box, a.dt, nve && a.limit, a.limit_val);
            pout[i] = p;
            a.image[i] = img;
            S.pos[threadIdx.x] = p; // this thread's next step, and the next step's walkers of this CTA (behind the barrier)
            S.img[threadIdx.x] = make_int4(img.x, img.y, img.z, im4.w);
            const float4 lp = S.lp[threadIdx.x];
            float3 dx = make_float3(__fsub_rn(p.x, __fmul_rn(S.lam[0], lp.x)), __fsub_rn(p.y, __fmul_rn(S.lam[1], lp.y)),
                                    __fsub_rn(p.z, __fmul_rn(S.lam[2], lp.z)));
            dx = min_image_exact_sel(dx, a.box, a.box.periodic.x, a.box.periodic.y, a.box.periodic.z);
            moved = dot_exact(dx) >= lp.w;
            }
        else if (ODD)
            pout[i] = pi4; // the chunk's final positions must be in buffer A
        a.vel[i] = v;
        S.vel[threadIdx.x] = v;
        }

    // ---- ticket + rebuild decision by the last CTA + grid-wide release ----
    const int any_moved = __syncthreads_or(moved ? 1 : 0);
    PERSIST_MARK(3)
    if (threadIdx.x == 0)
        {
        // release: this CTA's position stores (ordered before by the barrier above) before its ticket.  A release ATOMIC, not a
        // fence: the fence would also drop the SM's L1 lines under the CTAs that are still walking their neighbor rows
        const unsigned long long inc = 1ull | (any_moved ? (1ull << 32) : 0ull);
        unsigned long long old = 0ull;
#ifdef B200_FAULT_INJECT
        const bool lose = g_fault_lose_cta == blockIdx.x + 1u;
        if (lose) g_fault_lose_cta = 0u;
        if (!lose)
#endif
        asm volatile("atom.release.gpu.global.add.u64 %0, [%1], %2;" : "=l"(old) : "l"(persist_ticket(st)), "l"(inc) : "memory");
        PERSIST_MARK(4)
        const bool last_cta = (unsigned int)old == gridDim.x - 1u;
        if (last_cta)
            {
            // The LAST CTA takes the rebuild decision on a register copy of the control block (six independent 16-byte loads,
            // one round trip) instead of walking it with a dozen dependent loads while 999 CTAs spin: that serial section was
            // 4.7 us of a 17 us step (profiles/r2y_trace_*).  Nobody else writes the block while a launch runs.
            StepState loc;
            uint4* lq = reinterpret_cast<uint4*>(&loc);
            static_assert(sizeof(StepState) == 96, "StepState is copied as six 16-byte words");
#pragma unroll
            for (int q = 0; q < 6; q++)
                asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                             : "=r"(lq[q].x), "=r"(lq[q].y), "=r"(lq[q].z), "=r"(lq[q].w)
                             : "l"(reinterpret_cast<const uint4*>(st) + q)
                             : "memory");
            loc.moved = ((old + inc) >> 32) ? 1u : 0u;
            loc.steps_left -= 1u;
            if (!last)
                decide_body<false>(&loc);
            else
                {
                loc.need_rebuild = 0u;
                loc.moved = 0u;
                }
            loc.epoch = epoch + 1u;
            const unsigned long long w = persist_pack_ctl(&loc, epoch + 1u);
            uint4* gq = reinterpret_cast<uint4*>(st);
            gq[0] = lq[0]; // timestep, ts_step_two, moved, need_rebuild
            gq[1] = lq[1]; // force_update, last_updated, every, r_buff_tiny
            gq[2] = lq[2]; // num_updates, num_forced, overflow, ntypes
            gq[3] = lq[3]; // steps_left, epoch, (the one-step kernel's ticket: untouched, zero)
            st->box_scaled = loc.box_scaled;
            *persist_ticket(st) = 0ull;
            persist_fence(); // release: the block and the ticket reset before any copy of the release word
#pragma unroll
            for (unsigned int c = 0; c < STEP_SYNC_COPIES; c++)
                asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(persist_release_word(st, c)), "l"(w) : "memory");
            }
        else if (!a.persist_strict)
            persist_fence(); // (experiment) the acquire side's L1 invalidation taken during the wait, see below
        unsigned int spins = 0;
        bool lost = false;
        unsigned long long w;
        const unsigned int want = (epoch + 1u) & 0xffffffu;
        const unsigned long long* word = persist_release_word(st, blockIdx.x);
        while (true)
            {
            w = ld_relaxed_u64(word);
            if ((unsigned int)(w >> 40) == want) break;
            if (++spins > (1u << 22)) // seconds: a CTA of this grid never arrived
                {
                lost = true;
                break;
                }
            // a poll comes back in ~50 ns, so a thousand spinning threads put ~20 G loads/s on the 16 release lines and 14 % of
            // the kernel's issued instructions are this loop (profiles/r2z_mdsteps_64k_lines.txt) -- yet a pause between polls
            // buys nothing: 13.24 / 13.18 / 13.34 / 13.67 us per step with 0 / 100 / 300 / 1000 ns (profiles/r2y_*sleep*).  Off.
            if (a.persist_sleep_ns) __nanosleep(a.persist_sleep_ns);
            }
        PERSIST_MARK(5)
        // acquire: the other CTAs' position stores of this step are visible to this SM from here on.
        // HOOMD_B200_PERSIST_EARLY_FENCE=1 (experiment, off) takes this fence BEFORE the spin instead: the positions are double
        // buffered, during a step the SM's CTAs read buffer X only, L1 is filled by demand loads alone and stores do not
        // allocate, so the only stale sectors of the buffer read next are those of the step before, and every CTA's fence
        // behind its own last read drops them -- the MEMBAR + CCTL.IVALL then runs while the CTA would spin anyway.  Same
        // bits in every test, but it leans on how the hardware fills L1 rather than on the PTX memory model (which wants the
        // fence behind the observing load), and it measured only 13.18 vs 13.38 us per step (profiles/r2y_*): not the default.
        if (a.persist_strict) persist_fence();
        persist_unpack_ctl(w, s_ctl);
        PERSIST_MARK(6)
        PERSIST_MARK_END()
        if (lost)
            {
            st->nvt_error = 3u;
            s_ctl[3] |= PERSIST_STOP_ERROR;
            }
        }
    epoch += 1u;
    __syncthreads();
    unsigned int stop = s_ctl[3];
    if (s_ctl[0] == 0u) stop |= PERSIST_STOP_DONE;
    if (last) stop |= PERSIST_STOP_DONE;
    return stop;
    }

template<class Evaluator, int TPP, int SHIFT_MODE>
__global__ void __launch_bounds__(MD_BLOCK, 7) k_md_steps(const __grid_constant__ MdStepArgs<Evaluator> a)
    {
    constexpr int PPB = MD_BLOCK / TPP;
    extern __shared__ unsigned char smem_raw[];
    __shared__ float4 s_force[PPB];
    __shared__ PersistShared<PPB> S;
    __shared__ unsigned int s_ctl[4];
    __shared__ unsigned int s_epoch;
    StepState* st = a.st;
    if (threadIdx.x == 0)
        {
        persist_read_ctl(st, s_ctl);
        // the epoch counts on from what the copies of the release word hold (all equal between launches)
        s_epoch = (unsigned int)(ld_relaxed_u64(persist_release_word(st, blockIdx.x)) >> 40);
        }
    __syncthreads();
    // grid-uniform: the control block only changes after EVERY CTA has taken its first ticket
    if (s_ctl[0] == 0u || (s_ctl[3] & PERSIST_STOP_ERROR)) return;
    unsigned int epoch = s_epoch;
    PairTables<Evaluator> T;
    pair_tables_load<Evaluator, SHIFT_MODE, false>(T, smem_raw, a.params, a.rcutsq, a.ronsq, a.ntypes);
    PersistRegs R;
    persist_init<Evaluator, TPP>(a, R, S, (a.N + gridDim.x - 1u) / gridDim.x, (s_ctl[2] & 2u) != 0u);
    while (true)
        {
        const bool was_last = s_ctl[0] == 1u;
        unsigned int stop = md_persist_step<Evaluator, TPP, SHIFT_MODE, false>(a, st, T, s_force, S, s_ctl, epoch, R);
        if (stop)
            {
            // stopped after an A -> B step that moved the particles: put this CTA's particles back into A (every CTA's reads of
            // A for that step are behind the barrier)
            const unsigned int ppb = (a.N + gridDim.x - 1u) / gridDim.x;
            const unsigned int i = blockIdx.x * ppb + threadIdx.x;
            if (!was_last && threadIdx.x < ppb && i < a.N) const_cast<float4*>(a.pos_in)[i] = S.pos[threadIdx.x];
            break;
            }
        stop = md_persist_step<Evaluator, TPP, SHIFT_MODE, true>(a, st, T, s_force, S, s_ctl, epoch, R);
        if (stop) break;
        }
#if (CUDART_VERSION >= 12040)
    // arm the IF node of the graph unit behind this one (its rebuild bracket); s_ctl holds the control block as of the exit
    if (a.use_cond && blockIdx.x == 0 && threadIdx.x == 0)
        cudaGraphSetConditional(a.handle, (s_ctl[0] != 0u && s_ctl[3] == PERSIST_STOP_REBUILD) ? 1u : 0u);
#endif
    }

// co-resident CTAs of k_md_steps on the current device (0: cooperative launch unavailable)
template<class Evaluator, int TPP, int SHIFT_MODE> static unsigned int md_steps_max_ctas()
    {
    static int cached = -1;
    if (cached < 0)
        {
        cached = 0;
        int dev = 0, coop = 0, sms = 0, per_sm = 0;
        if (cudaGetDevice(&dev) == cudaSuccess
            && cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev) == cudaSuccess && coop
            && cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess
            && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_md_steps<Evaluator, TPP, SHIFT_MODE>, MD_BLOCK, 0)
                   == cudaSuccess)
            cached = sms * per_sm;
        }
    return (unsigned int)cached;
    }

static unsigned int md_steps_sm_count()
    {
    static int sms = -1;
    if (sms < 0)
        {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess
            || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
            sms = 1;
        }
    return (unsigned int)sms;
    }

template<class Evaluator, int TPP, int SHIFT_MODE>
static cudaError_t md_steps_launch(const MdStepArgs<Evaluator>& a, bool probe_only, cudaStream_t stream)
    {
    const unsigned int per_block = MD_BLOCK / TPP;
    unsigned int grid = (a.N + per_block - 1) / per_block;
    const unsigned int max_ctas = md_steps_max_ctas<Evaluator, TPP, SHIFT_MODE>();
    if (grid == 0 || grid > max_ctas) return cudaErrorNotSupported;
    // HOOMD_B200_PERSIST_EVEN=1: a grid that is a multiple of the SM count, the particles spread evenly over it (62 instead of
    // 64 per CTA at 64k, 7 CTAs on every SM instead of 7 on some and 6 on others).  Measured no better: 35.2k vs 35.5k TPS,
    // 17.25 vs 17.0 us per step (profiles/r2r_*) -- the ragged CTAs cost what the balance gains.  Off by default.
    const unsigned int sms = md_steps_sm_count();
    const unsigned int even = (grid + sms - 1u) / sms * sms;
    static const bool spread = getenv("HOOMD_B200_PERSIST_EVEN") && atoi(getenv("HOOMD_B200_PERSIST_EVEN")) != 0;
    if (spread && even <= max_ctas) grid = even;
    if (probe_only) return cudaSuccess;
    void* args[] = {(void*)&a};
    return cudaLaunchCooperativeKernel((const void*)k_md_steps<Evaluator, TPP, SHIFT_MODE>, dim3(grid), dim3(MD_BLOCK), args, 0,
                                       stream);
    }

// cudaErrorNotSupported: this system / configuration keeps the one-step kernel
template<class Evaluator>
static cudaError_t launch_md_steps_t(const b200_pair_args& pa, const void* d_params, const FusedStepArgs& fa,
                                     const float4* pos_a, float4* pos_b, StepState* st, bool probe_only,
                                     unsigned long long next_handle, int use_cond, cudaStream_t stream)
    {
    if (fa.N == 0 || !box_is_orthorhombic(pa.box)) return cudaErrorNotSupported;
    const bool extras = fa.group_flag || fa.tally_acc || fa.use_lambda || fa.mg || fa.gs.tab || fa.gs.cta_order || fa.d_N;
    if (extras || pa.ntypes != 1 || pa.compute_virial || pa.shift_mode == B200_SHIFT_XPLOR || fa.extra_force
        || (fa.method != 1 && fa.method != 2))
        return cudaErrorNotSupported;
    MdStepArgs<Evaluator> a;
    fill_md_step_args(a, pa, (const float*)d_params, fa, pos_a, pos_b, 0, st, next_handle, use_cond);
    static const bool early = getenv("HOOMD_B200_PERSIST_EARLY_FENCE") && atoi(getenv("HOOMD_B200_PERSIST_EARLY_FENCE")) != 0;
    a.persist_strict = early ? 0 : 1;
    static const int sleep_ns = getenv("HOOMD_B200_PERSIST_SLEEP") ? atoi(getenv("HOOMD_B200_PERSIST_SLEEP")) : 0;
    a.persist_sleep_ns = (unsigned int)std::max(0, sleep_ns);
    const unsigned int tpp = pa.threads_per_particle ? pa.threads_per_particle : pick_pair_tpp(pa.N);
    const bool shift = pa.shift_mode == B200_SHIFT_SHIFT;
    switch (tpp)
        {
        case 2:
            return shift ? md_steps_launch<Evaluator, 2, B200_SHIFT_SHIFT>(a, probe_only, stream)
                         : md_steps_launch<Evaluator, 2, B200_SHIFT_NONE>(a, probe_only, stream);
        case 4:
            return shift ? md_steps_launch<Evaluator, 4, B200_SHIFT_SHIFT>(a, probe_only, stream)
                         : md_steps_launch<Evaluator, 4, B200_SHIFT_NONE>(a, probe_only, stream);
        case 8:
            return shift ? md_steps_launch<Evaluator, 8, B200_SHIFT_SHIFT>(a, probe_only, stream)
                         : md_steps_launch<Evaluator, 8, B200_SHIFT_NONE>(a, probe_only, stream);
        default:
            return cudaErrorNotSupported;
        }
    }
} // namespace b200
