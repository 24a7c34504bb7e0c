// Multi-GPU half of the engine (included by engine.cu): DomainDecomposition + Communicator driven from the device.
// Reference semantics: hoomd/DomainDecomposition.cc:323-372 (grid), :531-557 (local box: non-periodic along split axes),
// hoomd/Communicator.cc:1300-1366 (communicate(): migrate + exchange ghosts on rebuild steps, update ghosts otherwise),
// hoomd/md/NeighborListGPU.cc:117-122 (global vote on the distance check).
namespace b200
{
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

static void mg_check_errors(b200_system* s)
    {
    CUDA_OK(cudaMemcpyAsync(s->h_ms, s->d_ms.p, sizeof(MgState), cudaMemcpyDeviceToHost, s->stream));
    CUDA_OK(cudaStreamSynchronize(s->stream));
    const unsigned int e = s->h_ms->error;
    if (!e) return;
    std::string msg = "comm:";
    if (e & 1u) msg += " migration mailbox overflow (too many particles left the domain in one rebuild);";
    if (e & 2u) msg += " ghost segment overflow (raise capacity_factor);";
    if (e & 4u) msg += " local particle capacity exceeded (raise Ncap);";
    if (e & 8u) msg += " a peer rank did not reach the barrier within 4 s;";
    throw std::runtime_error(msg);
    }

// geometry, capacities and the slab; peers are attached later by b200_sys_mg_connect
static void mg_setup(b200_system* s, const b200_mg_params& p)
    {
    if (p.world < 1 || p.world > (unsigned int)MG_MAX_RANKS) throw std::runtime_error("comm: 1..8 ranks supported");
    if (p.grid[0] * p.grid[1] * p.grid[2] != p.world) throw std::runtime_error("comm: grid does not match the world size");
    if (!box_is_orthorhombic(p.global_box)) throw std::runtime_error("comm: orthorhombic boxes only");
    s->mg = true;
    s->global_box = p.global_box;
    s->r_ghost = p.r_ghost;
    MgConst& c = s->mgc;
    memset(&c, 0, sizeof(c));
    c.world = p.world;
    c.rank = p.rank;
    unsigned int r = p.rank;
    for (int a = 0; a < 3; a++)
        {
        s->mg_grid[a] = p.grid[a];
        s->mg_pos[a] = r % p.grid[a];
        r /= p.grid[a];
        }
    // local box: uniform cut planes (DomainDecomposition ctor with uniform fractions), non-periodic along split axes
    float lo[3], hi[3];
    unsigned char per[3];
    for (int a = 0; a < 3; a++)
        {
        const float L = p.global_box.L[a], glo = p.global_box.lo[a];
        c.split[a] = p.grid[a] > 1;
        c.Lg[a] = L;
        c.glo[a] = glo;
        c.ghi[a] = p.global_box.hi[a];
        if (c.split[a])
            {
            lo[a] = glo + L * ((float)s->mg_pos[a] / (float)p.grid[a]);
            hi[a] = (s->mg_pos[a] + 1 == p.grid[a]) ? p.global_box.hi[a]
                                                    : glo + L * ((float)(s->mg_pos[a] + 1) / (float)p.grid[a]);
            per[a] = 0;
            if (hi[a] - lo[a] < 2.0f * p.r_ghost)
                throw std::runtime_error("comm: domain narrower than two ghost layers");
            }
        else
            {
            lo[a] = glo;
            hi[a] = p.global_box.hi[a];
            per[a] = p.global_box.periodic[a];
            }
        c.lo[a] = lo[a];
        c.hi[a] = hi[a];
        }
    b200_box_init_lohi(&s->box, lo, hi, per);
    c.r_ghost = p.r_ghost;
    // neighbors, periodic shifts, ghost segments (identical on every rank: uniform domains)
    const double rho = (double)p.N_global
                       / ((double)p.global_box.L[0] * p.global_box.L[1] * p.global_box.L[2]);
    const double cf = p.capacity_factor > 0.0f ? p.capacity_factor : 1.5;
    unsigned int off = 0;
    for (int d = 0; d < MG_NDIR; d++)
        {
        const int dd[3] = {d % 3 - 1, (d / 3) % 3 - 1, d / 9 - 1};
        bool used = d != MG_DIR_STAY;
        int np[3];
        double vol = 1.0;
        for (int a = 0; a < 3; a++)
            {
            if (dd[a] != 0 && !c.split[a]) used = false;
            int q = (int)s->mg_pos[a] + dd[a];
            c.shift[d][a] = 0.0f;
            if (q < 0) { q += (int)p.grid[a]; c.shift[d][a] = c.Lg[a]; }
            if (q >= (int)p.grid[a]) { q -= (int)p.grid[a]; c.shift[d][a] = -c.Lg[a]; }
            np[a] = q;
            vol *= dd[a] != 0 ? (double)p.r_ghost : (double)(hi[a] - lo[a]);
            }
        c.dir_rank[d] = used ? (int)((np[2] * p.grid[1] + np[1]) * p.grid[0] + np[0]) : -1;
        c.seg_off[d] = off;
        c.seg_cap[d] = used ? (unsigned int)(cf * rho * vol) + 1024u : 0u;
        off += c.seg_cap[d];
        if (!used)
            for (int a = 0; a < 3; a++) c.shift[d][a] = 0.0f;
        }
    s->Gcap = off;
    c.Gcap = off;
    // receiver side: segment d is filled by the neighbor this rank lies in direction d of; its particles need +-L to sit next
    // to the local box when that step crossed the global periodic boundary (dd = -1 into the last domain, +1 into the first)
    std::vector<float> rs(MG_NDIR * 3, 0.0f);
    for (int d = 0; d < MG_NDIR; d++)
        {
        if (c.dir_rank[d] < 0) continue;
        const int dd[3] = {d % 3 - 1, (d / 3) % 3 - 1, d / 9 - 1};
        for (int a = 0; a < 3; a++)
            {
            if (dd[a] == -1 && s->mg_pos[a] + 1 == p.grid[a]) rs[3 * d + a] = c.Lg[a];
            if (dd[a] == 1 && s->mg_pos[a] == 0) rs[3 * d + a] = -c.Lg[a];
            }
        }
    s->recv_shift.alloc(MG_NDIR * 3);
    s->recv_shift.upload(rs.data(), rs.size(), s->stream);
    const unsigned int expect = p.N_global / p.world + 1;
    s->Ncap = p.Ncap ? p.Ncap : (unsigned int)(1.25 * expect) + 8192u;
    if (p.N_local > s->Ncap) throw std::runtime_error("comm: more local particles than Ncap");
    c.Ncap = s->Ncap;
    c.mig_cap = std::max(4096u, s->Ncap / 32u);
    s->N = p.N_local;
    // slab
    const size_t ntot = (size_t)s->Ncap + s->Gcap;
    size_t o = 0;
    s->off_posA = o; o = align_up(o + ntot * sizeof(float4), 256);
    s->off_posB = o; o = align_up(o + ntot * sizeof(float4), 256);
    s->off_tag = o; o = align_up(o + ntot * sizeof(unsigned int), 256);
    s->off_mig = o; o = align_up(o + (size_t)MG_NDIR * c.mig_cap * sizeof(MigElem), 256);
    s->off_migcnt = o; o = align_up(o + MG_NDIR * sizeof(unsigned int), 256);
    // flag words, then as many value words (the all-reduce that rides on a barrier, comm.cuh:mg_warp_barrier)
    s->off_flags = o; o = align_up(o + (size_t)2 * MG_NBARRIER * 2 * MG_MAX_RANKS * sizeof(unsigned long long), 256);
    s->slab_bytes = o;
    CUDA_OK(cudaMalloc(&s->slab, s->slab_bytes));
    CUDA_OK(cudaMemset(s->slab, 0, s->slab_bytes));
    CUDA_OK(cudaDeviceSynchronize());
    char* base = (char*)s->slab;
    s->pos.view((float4*)(base + s->off_posA), ntot);
    s->pos_alt.view((float4*)(base + s->off_posB), ntot);
    s->tag.view((unsigned int*)(base + s->off_tag), ntot);
    c.my_mig = (MigElem*)(base + s->off_mig);
    c.my_mig_cnt = (unsigned int*)(base + s->off_migcnt);
    // rank-local scratch
    s->d_ms.alloc(1);
    CUDA_OK(cudaHostAlloc((void**)&s->h_ms, sizeof(MgState), cudaHostAllocDefault));
    s->holes.alloc(s->Ncap); s->low_holes.alloc(s->Ncap); s->hole_mark.alloc(s->Ncap);
    s->plan_src.alloc(std::max(1u, s->Gcap));
    s->slot_dir.alloc(std::max(1u, s->Gcap));
    std::vector<unsigned char> sd(std::max(1u, s->Gcap), 0);
    for (int d = 0; d < MG_NDIR; d++)
        for (unsigned int k = 0; k < c.seg_cap[d]; k++) sd[c.seg_off[d] + k] = (unsigned char)d;
    s->slot_dir.upload(sd.data(), sd.size(), s->stream);
    s->pplan.alloc(7 * (size_t)s->Ncap);
    s->pplan_n.alloc(s->Ncap);
    s->gs_table.alloc(1);
    // chunk size of k_md_step = particles per CTA: fixed for the life of the system (it shapes cta_order)
    // (2 or 8 lanes per particle: the widths every pair evaluator is instantiated for)
    s->mg_n_global = p.N_global;
    s->mg_tpp = pick_pair_tpp(std::max(1u, p.N_global / p.world)) <= 2u ? 2u : 8u;
    const unsigned int ppb = md_step_block() / s->mg_tpp;
    const unsigned int nchunks = (s->Ncap + ppb - 1) / ppb;
    s->chunk_flag.alloc(nchunks);
    s->cta_order.alloc(nchunks);
    std::vector<unsigned int> ident(nchunks);
    for (unsigned int k = 0; k < nchunks; k++) ident[k] = k;
    s->cta_order.upload(ident.data(), nchunks, s->stream);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    }

static void mg_bind_local(b200_system* s)
    {
    MgLocal& l = s->mgl;
    l.ms = s->d_ms.p;
    l.st = s->d_state.p;
    l.vel = s->vel.p;
    l.accel = (float*)s->accel.p;
    l.image = (int*)s->image.p;
    l.tag = s->tag.p;
    l.holes = s->holes.p;
    l.low_holes = s->low_holes.p;
    l.hole_mark = s->hole_mark.p;
    l.plan_src = s->plan_src.p;
    l.slot_dir = s->slot_dir.p;
    l.pplan = s->pplan.p;
    l.pplan_n = s->pplan_n.p;
    l.chunk_flag = s->chunk_flag.p;
    l.cta_order = s->cta_order.p;
    l.ppb = md_step_block() / s->mg_tpp;
    l.nchunks = (unsigned int)s->cta_order.n;
    }

static void mg_connect(b200_system* s, const unsigned char* handles)
    {
    MgConst& c = s->mgc;
    for (unsigned int q = 0; q < c.world; q++)
        {
        if (q == c.rank)
            s->peer_base[q] = s->slab;
        else
            {
            cudaIpcMemHandle_t h;
            memcpy(&h, handles + (size_t)q * sizeof(cudaIpcMemHandle_t), sizeof(h));
            CUDA_OK(cudaIpcOpenMemHandle(&s->peer_base[q], h, cudaIpcMemLazyEnablePeerAccess));
            }
        c.flags[q] = (unsigned long long*)((char*)s->peer_base[q] + s->off_flags);
        }
    for (int d = 0; d < MG_NDIR; d++)
        {
        if (c.dir_rank[d] < 0) continue;
        char* pb = (char*)s->peer_base[c.dir_rank[d]];
        c.peer_posA[d] = (float4*)(pb + s->off_posA);
        c.peer_posB[d] = (float4*)(pb + s->off_posB);
        c.peer_tag[d] = (unsigned int*)(pb + s->off_tag);
        c.peer_mig[d] = (MigElem*)(pb + s->off_mig) + (size_t)d * c.mig_cap;
        c.peer_mig_cnt[d] = (unsigned int*)(pb + s->off_migcnt) + d;
        }
    // device copy of what the fused ghost update needs per direction
    GhostSendTable t;
    memset(&t, 0, sizeof(t));
    for (int d = 0; d < MG_NDIR; d++)
        {
        for (int k = 0; k < 3; k++) t.shift[d][k] = c.shift[d][k];
        t.peer_posA[d] = c.peer_posA[d];
        t.peer_posB[d] = c.peer_posB[d];
        }
    t.world = c.world;
    t.rank = c.rank;
    for (unsigned int q = 0; q < c.world; q++) t.flags[q] = c.flags[q];
    t.ms = s->d_ms.p;
    s->gs_table.upload(&t, 1, s->stream);
    CUDA_OK(cudaStreamSynchronize(s->stream));
    mg_bind_local(s);
    s->mg_connected = true;
    invalidate_graph(s);
    }

// the rebuild bracket of one rank: migrate -> ghost plan + first ghost write -> cell list -> neighbor list
// bonded forces / exclusions: tag -> index over locals and ghosts, then NeighborList::filterNlist on the fresh rows
static bool mg_has_tag_tables(const b200_system* s)
    {
    return s->bond_n.n != 0 || (s->exclusions_set && s->ex_max);
    }
static void mg_enqueue_tag_maps(b200_system* s, const unsigned int* enable, const float4* pos)
    {
    if (!mg_has_tag_tables(s)) return;
    CUDA_OK(launch_mg_rtag(s->mgc, s->mgl, pos, s->rtag.p, s->mg_n_global, enable, s->stream));
    enqueue_nlist_filter(s, enable);
    }

static void mg_enqueue_rebuild(b200_system* s, const unsigned int* enable, float4* pos, int buffer)
    {
    CUDA_OK(launch_mg_migrate(s->mgc, s->mgl, pos, enable, s->stream));
    CUDA_OK(launch_mg_ghost_plan(s->mgc, s->mgl, pos, buffer, enable, s->stream));
    b200_cell_list_args ca;
    cell_list_args(s, ca);
    ca.d_pos = (const float*)pos;
    CUDA_OK(launch_cell_list(&ca, enable, s->stream, &s->d_state.p->N));
    b200_nlist_args na;
    nlist_args(s, na);
    na.d_pos = (const float*)pos;
    OverflowHook hook; // every rank must keep stepping to the barrier: raise the flag only
    hook.overflow = &s->d_state.p->overflow;
    CUDA_OK(launch_nlist_binned(&na, enable, s->stream, hook));
    mg_enqueue_tag_maps(s, enable, pos);
    CUDA_OK(launch_mg_rebuild_vote(s->mgc, s->mgl, buffer, enable, s->stream));
    }
constexpr unsigned long long MG_REBUILD_KERNELS = 8 + 4 + 4 + 2; // migrate, ghost plan + CTA order, cell list, nlist + vote

static void mg_upload_state(b200_system* s, unsigned int steps_left)
    {
    StepState& st = *s->h_state;
    memset(&st, 0, sizeof(st));
    st.timestep = s->timestep;
    st.force_update = s->force_update ? 1u : 0u;
    st.last_updated = s->last_updated_tstep;
    st.every = s->every;
    st.r_buff_tiny = (s->r_buff < 1e-6f) ? 1u : 0u;
    st.ntypes = s->ntypes;
    st.steps_left = steps_left;
    st.N = s->N;
    CUDA_OK(cudaMemcpyAsync(s->d_state.p, s->h_state, sizeof(StepState), cudaMemcpyHostToDevice, s->stream));
    }

static void mg_download_state(b200_system* s)
    {
    CUDA_OK(cudaMemcpyAsync(s->h_state, s->d_state.p, sizeof(StepState), cudaMemcpyDeviceToHost, s->stream));
    CUDA_OK(cudaStreamSynchronize(s->stream));
    s->N = s->h_state->N;
    }

// prepRun under domain decomposition: a forced, stream-launched rebuild (every rank runs the same sequence, so the
// barriers inside pair up), Nmax growth as in NeighborList::compute, then the pair force on the local particles
static void mg_prep_run(b200_system* s, unsigned int ts)
    {
    if (!s->mg_connected) throw std::runtime_error("comm: peers not attached (b200_sys_mg_connect)");
    if (s->force_update || s->nl_first_compute)
        {
        nlist_build_head_list(s);
        mg_upload_state(s, 0);
        CUDA_OK(launch_mg_migrate(s->mgc, s->mgl, s->pos.p, nullptr, s->stream));
        CUDA_OK(launch_mg_ghost_plan(s->mgc, s->mgl, s->pos.p, 0, nullptr, s->stream));
        mg_download_state(s);
        mg_check_errors(s);
        bool overflowed;
        do
            {
            b200_cell_list_args ca;
            cell_list_args(s, ca);
            CUDA_OK(launch_cell_list(&ca, nullptr, s->stream, &s->d_state.p->N));
            cell_list_check_conditions(s);
            b200_nlist_args na;
            nlist_args(s, na);
            CUDA_OK(launch_nlist_binned(&na, nullptr, s->stream));
            overflowed = nlist_check_conditions(s);
            if (overflowed)
                {
                // head room: a later overflow inside a device-driven chunk costs a collective abort + resume
                for (unsigned int t = 0; t < s->ntypes; t++) s->h_Nmax[t] = ((s->h_Nmax[t] * 9u / 8u) + 7u) & ~7u;
                s->d_Nmax.upload(s->h_Nmax.data(), s->ntypes, s->stream);
                CUDA_OK(cudaStreamSynchronize(s->stream));
                invalidate_graph(s);
                s->num_overflows += 1;
                nlist_build_head_list(s);
                CUDA_OK(cudaMemsetAsync(s->nl_conditions.p, 0, sizeof(unsigned int) * s->ntypes, s->stream));
                }
            } while (overflowed);
        mg_enqueue_tag_maps(s, nullptr, s->pos.p);
        s->force_update = false;
        s->nl_first_compute = false;
        s->cl_first = false;
        s->cl_force = false;
        // NeighborList::needsUpdating counts a forced update only when the step has not been checked yet -- and
        // m_last_checked_tstep starts at 0, so the very first build at time step 0 is not counted (NeighborList.cc:1082-1091)
        if (s->last_checked_tstep != ts) s->num_forced += 1;
        s->last_updated_tstep = ts;
        s->kernels_last_run += MG_REBUILD_KERNELS + cell_list_kernels(s->dim.x * s->dim.y * s->dim.z, s->Ncap + s->Gcap) - 4u;
        }
    s->nl_last_computed = ts;
    s->last_checked_tstep = ts;
    b200_pair_args a;
    pair_args(s, a);
    CUDA_OK(b200_compute_pair_forces(s->pair_kind, &a, pair_params_ptr(s), s->stream));
    s->kernels_last_run += 1;
    if (s->bond_kind >= 0)
        {
        // second ForceCompute: net = pair + bond (Integrator::computeNetForce), as compute_net_force does on one GPU
        CUDA_OK(cudaMemcpyAsync(s->pair_force_copy.p, s->net_force.p, sizeof(float4) * s->Ncap, cudaMemcpyDeviceToDevice,
                                s->stream));
        if (s->compute_virial)
            CUDA_OK(cudaMemcpyAsync(s->pair_virial_copy.p, s->net_virial.p, sizeof(float) * 6 * (size_t)s->virial_pitch,
                                    cudaMemcpyDeviceToDevice, s->stream));
        enqueue_bond_forces(s, s->pos.p);
        CUDA_OK(launch_add_force(s->net_force.p, s->net_virial.p, s->virial_pitch, s->bond_force.p, s->bond_virial.p,
                                 s->virial_pitch, s->Ncap, s->compute_virial, s->stream));
        s->kernels_last_run += 2;
        check_bond_flags(s);
        }
    if (!s->accel_set)
        {
        CUDA_OK(launch_compute_accel((float*)s->accel.p, s->net_force.p, s->vel.p, s->N, s->stream));
        s->kernels_last_run += 1;
        s->accel_set = true;
        }
    }

// graph of two MD steps.  Default (ghost update + step barrier inside k_md_step, FusedStepArgs::mg == 2):
//     [arm h0 from need_rebuild] -> IF{rebuild from A} -> k_md_step(A->B) -> IF{rebuild from B} -> k_md_step(B->A)
// -- the single-GPU graph with the collective rebuild bracket in the IF bodies: the LAST CTA of k_md_step takes the all-ranks
// barrier, the vote and the rebuild decision and arms the next IF node.  HOOMD_B200_FUSED_GHOSTS=0 (round-1 sequence):
//     barrier -> IF{rebuild from A} -> k_md_step(A->B) -> ghosts(B) -> barrier -> IF{rebuild from B} -> k_md_step(B->A) -> ghosts(A)
static unsigned long long* mg_nvt_acc(b200_system* s) { return s->method == 3 ? s->nvt_acc.p : nullptr; }

static void mg_build_graph(b200_system* s)
    {
#if (CUDART_VERSION >= 12040)
    FusedStepArgs fa;
    fused_args(s, fa);
    const bool fused = fa.mg == 2;          // step barrier inside k_md_step
    const bool push = fa.gs.tab != nullptr; // ghost stores inside k_md_step
    const unsigned int* enable = &s->d_state.p->need_rebuild;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    bool capturing = false;
    try
        {
        CUDA_OK(cudaGraphCreate(&graph, 0));
        cudaGraphConditionalHandle h0, h1;
        CUDA_OK(cudaGraphConditionalHandleCreate(&h0, graph, 0, cudaGraphCondAssignDefault));
        CUDA_OK(cudaGraphConditionalHandleCreate(&h1, graph, 0, cudaGraphCondAssignDefault));
        CUDA_OK(cudaStreamBeginCaptureToGraph(s->stream, graph, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
        capturing = true;
        if (fused)
            CUDA_OK(launch_graph_pre(s->d_state.p, (unsigned long long)h0, s->stream));
        else
            CUDA_OK(launch_mg_step_barrier(s->mgc, s->mgl, (unsigned long long)h0, 1, s->stream, mg_nvt_acc(s)));
        cudaGraph_t body0 = add_if_node(s, graph, h0);
        enqueue_md_step(s, fa, 0, fused ? (unsigned long long)h1 : 0ull, fused ? 1 : 0);
        if (!push) CUDA_OK(launch_mg_ghost_send(s->mgc, s->mgl, s->pos_alt.p, 1, nullptr, s->stream));
        if (!fused) CUDA_OK(launch_mg_step_barrier(s->mgc, s->mgl, (unsigned long long)h1, 1, s->stream, mg_nvt_acc(s)));
        cudaGraph_t body1 = add_if_node(s, graph, h1);
        enqueue_md_step(s, fa, 1, 0ull, 0);
        if (!push) CUDA_OK(launch_mg_ghost_send(s->mgc, s->mgl, s->pos.p, 0, nullptr, s->stream));
        cudaGraph_t out = nullptr;
        CUDA_OK(cudaStreamEndCapture(s->stream, &out));
        capturing = false;
        CUDA_OK(cudaStreamBeginCaptureToGraph(s->stream, body0, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
        capturing = true;
        mg_enqueue_rebuild(s, enable, s->pos.p, 0);
        CUDA_OK(cudaStreamEndCapture(s->stream, &out));
        capturing = false;
        CUDA_OK(cudaStreamBeginCaptureToGraph(s->stream, body1, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
        capturing = true;
        mg_enqueue_rebuild(s, enable, s->pos_alt.p, 1);
        CUDA_OK(cudaStreamEndCapture(s->stream, &out));
        capturing = false;
        CUDA_OK(cudaGraphInstantiate(&exec, graph, 0));
        }
    catch (...)
        {
        if (capturing)
            {
            cudaGraph_t junk = nullptr;
            cudaStreamEndCapture(s->stream, &junk);
            }
        if (graph) cudaGraphDestroy(graph);
        cudaGetLastError();
        throw;
        }
    s->graph = graph;
    s->graph_exec = exec;
    s->graph_uses_cond = 1;
    s->graph_valid = true;
#else
    throw std::runtime_error("comm: conditional graph nodes need CUDA >= 12.4");
#endif
    }

static void mg_run_chunk(b200_system* s, unsigned int chunk)
    {
    unsigned int remaining = chunk;
    bool resume = false; // re-entering the step whose rebuild overflowed: its step one has already been taken
    while (remaining)
        {
        if (s->force_update) nlist_build_head_list(s);
        if (s->use_graph != 3 && !s->graph_valid)
            {
            invalidate_graph(s);
            mg_build_graph(s);
            }
        mg_upload_state(s, remaining);
        FusedStepArgs fa;
        fused_args(s, fa);
        // prologue: step one of the first update() + its ghost update.  Fused mode: one stand-alone barrier takes the first
        // vote and decision (every later one is taken by the last CTA of k_md_step); round-1 mode: the first barrier of the
        // graph does.
        const bool fused = fa.mg == 2;
        const bool push = fa.gs.tab != nullptr;
        if (!resume) CUDA_OK(launch_fused_step_one(fa, s->d_state.p, s->stream));
        if (s->method == 3)
            {
            // TwoStepNVTMTK: this rank's kinetic energy of the half-step velocities goes into the accumulators the first step
            // kernel reads; the first step barrier adds up the ranks (k_mg_step_barrier), the kernel advances the thermostat
            CUDA_OK(cudaMemsetAsync(s->nvt_acc.p, 0, sizeof(unsigned long long) * s->nvt_acc.n, s->stream));
            CUDA_OK(launch_nvt_seed(s->nvt_acc.p, s->nvt_partial.p, b200_system::NVT_BLOCKS, (const float*)s->vel.p, nullptr,
                                    s->N, s->nvt_scale, s->timestep & 1u, s->stream));
            s->kernels_last_run += 2;
            }
        CUDA_OK(launch_mg_ghost_send(s->mgc, s->mgl, s->pos.p, 0, nullptr, s->stream));
        if (fused) CUDA_OK(launch_mg_step_barrier(s->mgc, s->mgl, 0ull, 0, s->stream, mg_nvt_acc(s)));
        const unsigned int nreplay = (remaining + 1) / 2;
        if (s->use_graph == 3)
            {
            const unsigned int* enable = &s->d_state.p->need_rebuild;
            for (unsigned int k = 0; k < nreplay; k++)
                {
                if (!fused) CUDA_OK(launch_mg_step_barrier(s->mgc, s->mgl, 0ull, 0, s->stream, mg_nvt_acc(s)));
                mg_enqueue_rebuild(s, enable, s->pos.p, 0);
                enqueue_md_step(s, fa, 0, 0ull, 0);
                if (!push) CUDA_OK(launch_mg_ghost_send(s->mgc, s->mgl, s->pos_alt.p, 1, nullptr, s->stream));
                if (!fused) CUDA_OK(launch_mg_step_barrier(s->mgc, s->mgl, 0ull, 0, s->stream, mg_nvt_acc(s)));
                mg_enqueue_rebuild(s, enable, s->pos_alt.p, 1);
                enqueue_md_step(s, fa, 1, 0ull, 0);
                if (!push) CUDA_OK(launch_mg_ghost_send(s->mgc, s->mgl, s->pos.p, 0, nullptr, s->stream));
                }
            }
        else
            for (unsigned int k = 0; k < nreplay; k++) CUDA_OK(cudaGraphLaunch(s->graph_exec, s->stream));
        mg_download_state(s);
        const StepState st = *s->h_state;
        mg_check_errors(s);
        check_bond_flags(s);
        cell_list_check_conditions(s);
        const unsigned long long rebuilds = (unsigned long long)st.num_updates + st.num_forced;
        s->kernels_last_run += (fused ? 3ull : 2ull) + (fused ? 3ull : (push ? 4ull : 6ull)) * nreplay
                               + (MG_REBUILD_KERNELS + cell_list_kernels(s->dim.x * s->dim.y * s->dim.z, s->Ncap + s->Gcap) - 4u) * rebuilds;
        s->num_updates += st.num_updates;
        s->num_forced += st.num_forced;
        s->nl_first_compute = false;
        if (st.abort)
            {
            // collective Nmax overflow: the aborted step has its step one and its migration behind it but no force yet.
            // Grow Nmax where needed (NeighborList::checkConditions), then redo the step's rebuild as a forced update.
            if (st.abort == 2u)
                CUDA_OK(cudaMemcpyAsync(s->pos.p, s->pos_alt.p, sizeof(float4) * s->N, cudaMemcpyDeviceToDevice, s->stream));
            if (nlist_check_conditions(s))
                {
                for (unsigned int t = 0; t < s->ntypes; t++) s->h_Nmax[t] = ((s->h_Nmax[t] * 9u / 8u) + 7u) & ~7u;
                s->d_Nmax.upload(s->h_Nmax.data(), s->ntypes, s->stream);
                CUDA_OK(cudaStreamSynchronize(s->stream));
                invalidate_graph(s);
                }
            CUDA_OK(cudaMemsetAsync(s->nl_conditions.p, 0, sizeof(unsigned int) * s->ntypes, s->stream));
            s->num_overflows += 1;
            // the redo below is counted as one more forced update: the reference's overflow loop rebuilds inside ONE
            // NeighborList::compute and counts it once (hoomd/md/NeighborList.cc:226-244)
            s->num_forced -= 1;
            s->timestep = st.timestep - 1u; // the decision of the aborted step is taken again
            s->last_updated_tstep = st.last_updated;
            s->force_update = true;
            remaining = st.steps_left;
            resume = true;
            continue;
            }
        s->timestep = st.timestep;
        s->last_updated_tstep = st.last_updated;
        s->force_update = st.force_update != 0;
        s->last_checked_tstep = st.timestep;
        s->last_check_result = st.need_rebuild != 0;
        s->nl_last_computed = st.timestep;
        remaining = 0;
        }
    }
} // namespace b200
