// pf_app_host.inl - host side of the entry points around the stepping path: device-side initial
// conditions, overlapped snapshots, the reference host's output files, the headless colour-map view.
// Included at the end of pf_solver.cu (same translation unit: it uses struct pf_solver and the error macros).

extern "C" {

// The reference's reduction micro-benchmark (run_benchmarks, simulation.cu:1300-1358) on this library's reductions.
int pf_bench_reductions(pf_solver* s, size_t n, int value_dtype, int reps, double* seconds_max, double* seconds_stats, double* max_value)
{
    if(int rc = bind(s)) return rc;
    if(n == 0 || reps < 1 || (value_dtype != PF_F32 && value_dtype != PF_F64)) return fail(s, PF_ERR_INVALID, "pf_bench_reductions: bad arguments");
    const size_t es = pf_sizeof_real(value_dtype);
    void *a = nullptr, *b = nullptr;
    PF_CUDA(s, cudaMalloc(&a, n * es));
    if(cudaMalloc(&b, n * es) != cudaSuccess) { cudaFree(a); cudaGetLastError(); return fail(s, PF_ERR_OOM, "pf_bench_reductions: out of memory"); }
    const int fill_blocks = elementwise_blocks(s, n, 256);
    const int red_blocks = (int) std::max<size_t>(1, std::min<size_t>((n / (16 / es) + 255) / 256, (size_t) s->sm_count * 8));
    const int st_blocks = std::min<int>(CG_MAX_BLOCKS, elementwise_blocks(s, n, CG_THREADS));
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    std::vector<float> t_max, t_stats;
    int rc = PF_OK;
    auto ck = [&](cudaError_t e) { if(e != cudaSuccess && rc == PF_OK) rc = fail(s, PF_ERR_CUDA, "pf_bench_reductions: %s", cudaGetErrorString(e)); };
    ck(cudaEventCreate(&e0)); ck(cudaEventCreate(&e1));
    if(value_dtype == PF_F64) { k_fill_hash<double><<<fill_blocks, 256, 0, s->stream>>>(n, (double*) a, 1u); k_fill_hash<double><<<fill_blocks, 256, 0, s->stream>>>(n, (double*) b, 2u); }
    else                      { k_fill_hash<float><<<fill_blocks, 256, 0, s->stream>>>(n, (float*) a, 1u);  k_fill_hash<float><<<fill_blocks, 256, 0, s->stream>>>(n, (float*) b, 2u); }
    s->launches += 2;
    for(int r = -2; r < reps && rc == PF_OK; r++)                         // two warm-up rounds
    {
        float ms = 0;
        ck(cudaEventRecord(e0, s->stream));
        ck(cudaMemsetAsync(s->d_err, 0, sizeof(unsigned long long), s->stream));
        if(value_dtype == PF_F64) k_reduce_maxabs<double><<<red_blocks, 256, 0, s->stream>>>(n, (const double*) a, s->d_err);
        else                      k_reduce_maxabs<float><<<red_blocks, 256, 0, s->stream>>>(n, (const float*) a, s->d_err);
        ck(cudaEventRecord(e1, s->stream)); ck(cudaEventSynchronize(e1)); ck(cudaEventElapsedTime(&ms, e0, e1));
        if(r >= 0) t_max.push_back(ms);
        ck(cudaEventRecord(e0, s->stream));
        if(value_dtype == PF_F64) k_stats_delta<double><<<st_blocks, CG_THREADS, 0, s->stream>>>(n, (const double*) a, (const double*) b, s->d_stats);
        else                      k_stats_delta<float><<<st_blocks, CG_THREADS, 0, s->stream>>>(n, (const float*) a, (const float*) b, s->d_stats);
        ck(cudaEventRecord(e1, s->stream)); ck(cudaEventSynchronize(e1)); ck(cudaEventElapsedTime(&ms, e0, e1));
        if(r >= 0) t_stats.push_back(ms);
        s->launches += 2;
    }
    ck(cudaGetLastError());
    if(rc == PF_OK)
    {
        ck(cudaMemcpyAsync(s->h_err, s->d_err, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s->stream));
        ck(cudaStreamSynchronize(s->stream));
    }
    if(rc == PF_OK)
    {
        std::sort(t_max.begin(), t_max.end()); std::sort(t_stats.begin(), t_stats.end());
        if(seconds_max) *seconds_max = t_max[t_max.size() / 2] * 1e-3;
        if(seconds_stats) *seconds_stats = t_stats[t_stats.size() / 2] * 1e-3;
        if(max_value) *max_value = value_dtype == PF_F64 ? decode_err<double>(s->h_err[0]) : (double) decode_err<float>(s->h_err[0]);
    }
    if(e0) cudaEventDestroy(e0);
    if(e1) cudaEventDestroy(e1);
    cudaFree(a); cudaFree(b);
    return rc;
}

int pf_init_fields(pf_solver* s, const pf_params* params, const pf_initial* init, void* F, void* U)
{
    if(int rc = bind(s)) return rc;
    if(!params || !init || !F || !U) return fail(s, PF_ERR_INVALID, "pf_init_fields: NULL argument");
    if(params->nx != s->nx || params->ny != s->ny) return fail(s, PF_ERR_INVALID, "pf_init_fields: params %dx%d != solver %dx%d", params->nx, params->ny, s->nx, s->ny);
    InitArgs a;
    a.inside_phi = init->inside_phi; a.inside_T = init->inside_T; a.outside_phi = init->outside_phi; a.outside_T = init->outside_T;
    a.cx = init->circle_center[0]; a.cy = init->circle_center[1]; a.radius = init->circle_radius; a.fade = init->circle_fade;
    a.sq_x0 = init->square_from[0]; a.sq_y0 = init->square_from[1]; a.sq_x1 = init->square_to[0]; a.sq_y1 = init->square_to[1];
    a.L0 = params->L0; a.xi = params->xi;
    a.nx = s->nx; a.ny_global = s->ny;
    a.row0 = s->slab.y0 - s->ghost_lo; a.rows = s->rows_alloc;
    a.do_exact = params->do_exact != 0;
    a.wrap = params->Phi_boundary == PF_BOUNDARY_PERIODIC;
    dim3 block(256), grid((unsigned)((s->nx + 255) / 256), (unsigned) s->rows_alloc);
    if(s->dtype == PF_F64) k_init_fields<double><<<grid, block, 0, s->stream>>>(a, (double*) F, (double*) U);
    else                   k_init_fields<float><<<grid, block, 0, s->stream>>>(a, (float*) F, (float*) U);
    PF_LAUNCH_CHECK(s);
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    return PF_OK;
}

int pf_snapshot_begin(pf_solver* s, int slot, int n_fields, const void* const* dev_fields)
{
    if(int rc = bind(s)) return rc;
    if(slot < 0 || slot >= PF_SNAPSHOT_SLOTS) return fail(s, PF_ERR_INVALID, "pf_snapshot_begin: slot %d", slot);
    if(n_fields <= 0 || n_fields > PF_SNAPSHOT_MAX_FIELDS || !dev_fields) return fail(s, PF_ERR_INVALID, "pf_snapshot_begin: bad field list");
    for(int i = 0; i < n_fields; i++) if(!dev_fields[i]) return fail(s, PF_ERR_INVALID, "pf_snapshot_begin: field %d is NULL", i);
    pf_solver::SnapSlot& sl = s->snap[slot];
    const size_t owned = (size_t) s->nx * (size_t) s->slab.ny_local;
    const size_t bytes = owned * sizeof(double) * (size_t) n_fields;
    if(!s->snap_stream) PF_CUDA(s, cudaStreamCreateWithFlags(&s->snap_stream, cudaStreamNonBlocking));
    if(!sl.staged) PF_CUDA(s, cudaEventCreateWithFlags(&sl.staged, cudaEventDisableTiming));
    if(!sl.landed) PF_CUDA(s, cudaEventCreateWithFlags(&sl.landed, cudaEventDisableTiming));
    if(sl.pending) { PF_CUDA(s, cudaEventSynchronize(sl.landed)); sl.pending = false; }      // the slot's previous copy must have left the staging buffer
    if(sl.bytes < bytes)
    {
        if(sl.dev) { PF_CUDA(s, cudaFree(sl.dev)); sl.dev = nullptr; }
        if(sl.host) { PF_CUDA(s, cudaFreeHost(sl.host)); sl.host = nullptr; }
        sl.bytes = 0;
        PF_CUDA(s, cudaMalloc(&sl.dev, bytes));
        PF_CUDA(s, cudaMallocHost((void**) &sl.host, bytes));
        sl.bytes = bytes;
    }
    sl.n_fields = n_fields;
    const size_t skip = (size_t) s->ghost_lo * (size_t) s->nx;
    for(int i = 0; i < n_fields; i++)
    {
        double* dst = (double*) sl.dev + (size_t) i * owned;
        if(s->dtype == PF_F64)
            PF_CUDA(s, cudaMemcpyAsync(dst, (const double*) dev_fields[i] + skip, owned * sizeof(double), cudaMemcpyDeviceToDevice, s->stream));
        else
        {
            k_convert<double, float><<<elementwise_blocks(s, owned, 256), 256, 0, s->stream>>>(owned, (const float*) dev_fields[i] + skip, dst);
            PF_LAUNCH_CHECK(s);
        }
    }
    PF_CUDA(s, cudaEventRecord(sl.staged, s->stream));
    PF_CUDA(s, cudaStreamWaitEvent(s->snap_stream, sl.staged, 0));
    PF_CUDA(s, cudaMemcpyAsync(sl.host, sl.dev, bytes, cudaMemcpyDeviceToHost, s->snap_stream));
    PF_CUDA(s, cudaEventRecord(sl.landed, s->snap_stream));
    sl.pending = true;
    return PF_OK;
}

int pf_snapshot_wait(pf_solver* s, int slot, const double** host_fields)
{
    if(int rc = bind(s)) return rc;
    if(slot < 0 || slot >= PF_SNAPSHOT_SLOTS || !host_fields) return fail(s, PF_ERR_INVALID, "pf_snapshot_wait: bad argument");
    pf_solver::SnapSlot& sl = s->snap[slot];
    if(!sl.landed || sl.n_fields == 0) return fail(s, PF_ERR_INVALID, "pf_snapshot_wait: slot %d was never begun", slot);
    PF_CUDA(s, cudaEventSynchronize(sl.landed));
    sl.pending = false;
    const size_t owned = (size_t) s->nx * (size_t) s->slab.ny_local;
    for(int i = 0; i < sl.n_fields; i++) host_fields[i] = sl.host + (size_t) i * owned;
    return PF_OK;
}

int pf_render_rgba(pf_solver* s, const void* dev_field, float display_min, float display_max, uint8_t* host_rgba)
{
    if(int rc = bind(s)) return rc;
    if(!dev_field || !host_rgba) return fail(s, PF_ERR_INVALID, "pf_render_rgba: NULL argument");
    const size_t owned = (size_t) s->nx * (size_t) s->slab.ny_local, bytes = owned * 4;
    if(s->rgba_bytes < bytes)
    {
        if(s->d_rgba) { PF_CUDA(s, cudaStreamSynchronize(s->stream)); PF_CUDA(s, cudaFree(s->d_rgba)); s->d_rgba = nullptr; s->rgba_bytes = 0; }
        PF_CUDA(s, cudaMalloc(&s->d_rgba, bytes));
        s->rgba_bytes = bytes;
    }
    const size_t skip = (size_t) s->ghost_lo * (size_t) s->nx;
    const int blocks = elementwise_blocks(s, owned, 256);
    if(s->dtype == PF_F64) k_render_rgba<double><<<blocks, 256, 0, s->stream>>>(owned, (const double*) dev_field + skip, display_min, display_max, (uchar4*) s->d_rgba);
    else                   k_render_rgba<float><<<blocks, 256, 0, s->stream>>>(owned, (const float*) dev_field + skip, display_min, display_max, (uchar4*) s->d_rgba);
    PF_LAUNCH_CHECK(s);
    PF_CUDA(s, cudaMemcpyAsync(host_rgba, s->d_rgba, bytes, cudaMemcpyDeviceToHost, s->stream));
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    return PF_OK;
}

} // extern "C"

#include "pf_io.inl"
