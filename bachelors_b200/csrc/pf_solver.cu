// pf_solver.cu - host side of the C-ABI declared in include/pf_solver.h.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -std=c++17
//        -Xcompiler -fPIC,-ffp-contract=off -shared   (see __graft_entry__.build()).
// No CPU fallback exists: every entry point that touches fields needs a CUDA device.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include <algorithm>
#include <mutex>

#include "../../include/pf_solver.h"
#include "pf_common.cuh"
#include "pf_fused.cuh"
#include "pf_stream.cuh"
#include "pf_euler.cuh"
#include "pf_general.cuh"
#include "pf_si_coop.cuh"
#include "pf_app.cuh"
#include "pf_halo.cuh"

using namespace pf;

// -------------------------------------------------------------------------------------------------------
enum { PF_SCRATCH_SLOTS = 16 };

struct pf_solver {
    int device = 0;
    int dtype = PF_F64;
    int nx = 0, ny = 0;                 // global grid
    pf_slab slab = {};
    int ghost_lo = 0, ghost_hi = 0, rows_alloc = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;

    void* scratch[PF_SCRATCH_SLOTS] = {};   // workspace arena: whole fields, allocated on first use, freed in pf_destroy
    void* partials = nullptr;               // 4 x CG_MAX_BLOCKS reals (as doubles)
    CgState* d_cg = nullptr;                // [2]: Phi solve, T solve
    StatsPartial* d_stats = nullptr;        // CG_MAX_BLOCKS
    unsigned long long* d_err = nullptr;    // [2] error-norm bit patterns + (as unsigned) the arrival ticket in word [2]
    bool d_err_dirty = true;                // the words may be non-zero (the hand-over path needs them zero and leaves them zero)
    unsigned long long* h_mail = nullptr;   // [3] mapped pinned: {max|e_Phi| bits, max|e_T| bits, sequence number} written by the attempt kernel
    unsigned long long mail_seq = 0;
    void* staging = nullptr; size_t staging_bytes = 0;

    // pinned host mirrors
    unsigned long long* h_err = nullptr;    // [2]
    CgState* h_cg = nullptr;                // [2]
    StatsPartial* h_stats = nullptr;        // CG_MAX_BLOCKS

    // state the reference keeps in function statics
    double adaptive_tau = 0;                // simulation.cu:363
    int last_num_steps = 0;                 // simulation.cu:936

    unsigned stream_attr_done = 0;          // bit per streaming-kernel instance
    unsigned halo_attr_done = 0;            // same for k_stream_step_halo
    bool halo_in_kernel = false;            // the last fused launch did the halo hand-over itself
    int stream_resident[32];                // co-resident CTAs per SM of each streaming-kernel instance (0 = not queried yet)
    unsigned* run_progress = nullptr;       // multi-step launches: steps finished by each unit (+ 1 flag word), see StreamRun
    int run_progress_units = 0;
    unsigned run_base = 0;                  // progress value meaning "nothing of the next launch done yet"
    // pf_run: RUN_GRAPH_STEPS single-launch steps (ping-pong baked in) captured once and replayed as a CUDA graph
    cudaGraphExec_t run_graph = nullptr;
    struct RunGraphKey { const void *F, *U, *nF, *nU; pf_params p; } run_graph_key = {};
    bool run_graphs_ok = true;              // PF_OPT_RUN_GRAPHS; cleared when a capture fails
    int si_blocks_per_sm = 4;               // PF_OPT_SI_BLOCKS_PER_SM: blocks of the CG kernels per SM (1 .. 4)
    int run_steps = 1;                      // steps the next fused launch advances (set by pf_run around its launch)
    bool run_used = false;                  // the last fused launch really was a multi-step launch
    int coop_blocks_per_sm[4] = {-1, -1, -1, -1};   // co-resident CTAs per SM of k_si_step_coop<Real, ANISO> (-1: not queried yet)
    unsigned long long* ll_slots = nullptr; // flagged partial sums of the cooperative semi-implicit step (pf_si_coop.cuh GridSum)
    unsigned ll_epoch = 1;                  // next unused epoch (0 = the cleared slots)
    bool plain_grid_sync = false;           // PF_OPT_SI_GRID_SYNC: cooperative grid.sync() + partial arrays instead of the flagged sums
    bool force_per_kernel_cg = false;       // PF_OPT_SI_PER_KERNEL: take the per-kernel CG path
    int host_bands = 8;                     // PF_OPT_HOST_BANDS: row bands of the pipelined pf_step_host
    bool halo_in_kernel_ok = true;          // PF_OPT_HALO_IN_KERNEL: linked slabs hand their rows over inside the RK4 step kernel
    bool rkm_keep_errors = false;           // rows variant: do not zero the error words inside the attempt
    bool rkm_fused = true;                  // path chosen by rkm_begin for the current adaptive step
    bool rkm_handover = false;              // rk4_adaptive_step: the attempt kernel mails its norms to the host (see FusedArgs::err_host)
    int win_y0 = 0, win_y1 = 0;             // row window of the *_rows_async entry points (global rows); empty = all owned rows
    cudaStream_t aux_stream = nullptr;      // second stream: boundary rows of a slab run here while the interior runs on `stream`
    cudaStream_t launch_stream = nullptr;   // stream the next fused launch goes to (stream or aux_stream)
    cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;   // pf_step_host: copy streams (created on first use)
    enum { HOST_BANDS_MAX = 32 };
    cudaEvent_t ev_up[HOST_BANDS_MAX] = {}, ev_done[HOST_BANDS_MAX] = {};
    cudaEvent_t ev_host_begin = nullptr;
    // snapshots (pf_snapshot_*): device staging + pinned host buffer per slot, one copy stream
    struct SnapSlot { void* dev = nullptr; double* host = nullptr; size_t bytes = 0; int n_fields = 0; cudaEvent_t staged = nullptr, landed = nullptr; bool pending = false; };
    SnapSlot snap[PF_SNAPSHOT_SLOTS];
    cudaStream_t snap_stream = nullptr;
    void* d_rgba = nullptr; size_t rgba_bytes = 0;      // pf_render_rgba
    // pf_si_*: the semi-implicit step in phases (slabs); what pf_si_begin_async was given
    struct SiCtx { pf_params p; const void *F = nullptr, *U = nullptr, *U_base = nullptr; void *nF = nullptr, *nU = nullptr; bool active = false; } si;
    unsigned* halo_flags = nullptr;         // device-initiated halo exchange (pf_halo.cuh): words the neighbours write through peer mappings
    struct HaloLink {                       // pf_halo_link: the neighbours' mapped fields and flag blocks; every whole-slab step exchanges through them
        bool on = false;
        int n_fields = 0;
        void* local[8] = {}; void* peer[2][8] = {};
        size_t first_row[2] = {0, 0};
        unsigned* peer_flags[2] = {nullptr, nullptr};
        unsigned e = 0;                     // exchanges completed so far
    } halo;
    int64_t launches = 0;
    std::string err;

    size_t real_size() const { return dtype == PF_F64 ? 8 : 4; }
    size_t field_elems() const { return (size_t) nx * (size_t) rows_alloc; }
    size_t field_bytes() const { return field_elems() * real_size(); }
    bool is_full_grid() const { return slab.y0 == 0 && slab.ny_local == ny && !slab.has_lo && !slab.has_hi; }
};

static thread_local std::string g_create_error;
static int halo_wait_both(pf_solver* s);

static int fail(pf_solver* s, int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if(s) s->err = buf; else g_create_error = buf;
    return code;
}

#define PF_CUDA(s, call) do { cudaError_t e_ = (call); if(e_ != cudaSuccess) \
    return fail((s), e_ == cudaErrorMemoryAllocation ? PF_ERR_OOM : PF_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while(0)

#define PF_LAUNCH_CHECK(s) do { (s)->launches++; cudaError_t e_ = cudaGetLastError(); if(e_ != cudaSuccess) \
    return fail((s), PF_ERR_CUDA, "kernel launch failed: %s (%s:%d)", cudaGetErrorString(e_), __FILE__, __LINE__); } while(0)

static int bind(pf_solver* s)
{
    if(!s) return PF_ERR_INVALID;
    PF_CUDA(s, cudaSetDevice(s->device));
    return PF_OK;
}

static int get_scratch(pf_solver* s, int slot, void** out)
{
    if(slot < 0 || slot >= PF_SCRATCH_SLOTS) return fail(s, PF_ERR_INVALID, "bad scratch slot %d", slot);
    if(!s->scratch[slot])
    {
        PF_CUDA(s, cudaMalloc(&s->scratch[slot], s->field_bytes()));
        PF_CUDA(s, cudaMemsetAsync(s->scratch[slot], 0, s->field_bytes(), s->stream));
    }
    *out = s->scratch[slot];
    return PF_OK;
}

static inline int elementwise_blocks(const pf_solver* s, size_t N, int threads)
{
    size_t want = (N + threads - 1) / threads;
    size_t cap = (size_t) s->sm_count * 16;
    return (int) std::max<size_t>(1, std::min(want, cap));
}

// -------------------------------------------------------------------------------------------------------
extern "C" {

int pf_abi_version(void) { return PF_ABI_VERSION; }

int pf_device_count(void)
{
    int n = 0;
    if(cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char* pf_last_error(const pf_solver* s) { return s ? s->err.c_str() : g_create_error.c_str(); }

size_t pf_sizeof_real(int dtype) { return dtype == PF_F64 ? 8 : (dtype == PF_F32 ? 4 : 0); }

int pf_required_ghost(int solver_type)
{
    switch(solver_type)
    {
        case PF_SOLVER_EXPLICIT_EULER: return 1;
        case PF_SOLVER_EXPLICIT_RK4: return 4;
        case PF_SOLVER_EXPLICIT_RK4_ADAPTIVE: return 5;
        case PF_SOLVER_SEMI_IMPLICIT: return 1;
        default: return 0;
    }
}

int pf_create_slab(pf_solver** out, int device, int dtype, int nx, int ny_global, const pf_slab* slab)
{
    if(!out) return fail(nullptr, PF_ERR_INVALID, "pf_create: out is NULL");
    *out = nullptr;
    if(dtype != PF_F32 && dtype != PF_F64) return fail(nullptr, PF_ERR_INVALID, "pf_create: bad dtype %d", dtype);
    if(nx <= 0 || ny_global <= 0) return fail(nullptr, PF_ERR_INVALID, "pf_create: bad grid %dx%d", nx, ny_global);
    if((long long) nx * (long long) ny_global > 0x7fffffffLL) return fail(nullptr, PF_ERR_INVALID, "pf_create: nx*ny exceeds int32 (the reference's csize, cuda_util.cuh:19)");
    if(!slab || slab->ny_local <= 0 || slab->y0 < 0 || slab->y0 + slab->ny_local > ny_global || slab->ghost < 0)
        return fail(nullptr, PF_ERR_INVALID, "pf_create: bad slab");
    // a neighbour's ghost rows are filled from this slab's first / last `ghost` OWNED rows (pf_halo_region): thinner slabs
    // would need rows from two ranks away
    if((slab->has_lo || slab->has_hi) && slab->ny_local < slab->ghost)
        return fail(nullptr, PF_ERR_INVALID, "pf_create: a slab with a neighbour owns %d rows, fewer than its ghost depth %d", slab->ny_local, slab->ghost);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if(e != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(nullptr, PF_ERR_NO_DEVICE, "no CUDA device: %s (this library has no CPU fallback)", cudaGetErrorString(e)); }
    if(device < 0 || device >= ndev) return fail(nullptr, PF_ERR_INVALID, "pf_create: device %d out of range (%d devices)", device, ndev);

    pf_solver* s = new pf_solver();
    memset(s->stream_resident, 0, sizeof s->stream_resident);
    s->device = device; s->dtype = dtype; s->nx = nx; s->ny = ny_global; s->slab = *slab;
    s->ghost_lo = slab->has_lo ? slab->ghost : 0;
    s->ghost_hi = slab->has_hi ? slab->ghost : 0;
    s->rows_alloc = s->ghost_lo + slab->ny_local + s->ghost_hi;
    auto bail = [&](int code) { g_create_error = s->err; pf_destroy(s); return code; };
    #define PF_CREATE_CUDA(call) do { cudaError_t e_ = (call); if(e_ != cudaSuccess) { \
        fail(s, e_ == cudaErrorMemoryAllocation ? PF_ERR_OOM : PF_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); return bail(e_ == cudaErrorMemoryAllocation ? PF_ERR_OOM : PF_ERR_CUDA); } } while(0)
    PF_CREATE_CUDA(cudaSetDevice(device));
    PF_CREATE_CUDA(cudaDeviceGetAttribute(&s->sm_count, cudaDevAttrMultiProcessorCount, device));
    PF_CREATE_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    {
        // the auxiliary stream carries a slab's edge rows and their hand-over to the neighbours: its small kernels must get SM slots
        // ahead of the interior's CTAs
        int prio_lo = 0, prio_hi = 0;
        PF_CREATE_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        PF_CREATE_CUDA(cudaStreamCreateWithPriority(&s->aux_stream, cudaStreamNonBlocking, prio_hi));
    }
    PF_CREATE_CUDA(cudaMalloc(&s->partials, 4 * CG_MAX_BLOCKS * sizeof(double)));
    PF_CREATE_CUDA(cudaMalloc(&s->ll_slots, (4 * CG_MAX_BLOCKS + 4) * sizeof(unsigned long long)));   // GridSum: 2 sets x 2 words per block + 2 x 2 result words
    PF_CREATE_CUDA(cudaMemset(s->ll_slots, 0, (4 * CG_MAX_BLOCKS + 4) * sizeof(unsigned long long)));
    PF_CREATE_CUDA(cudaMalloc((void**) &s->d_cg, 2 * sizeof(CgState)));
    PF_CREATE_CUDA(cudaMalloc((void**) &s->d_stats, CG_MAX_BLOCKS * sizeof(StatsPartial)));
    PF_CREATE_CUDA(cudaMalloc((void**) &s->d_err, 3 * sizeof(unsigned long long)));
    PF_CREATE_CUDA(cudaMemset(s->d_err, 0, 3 * sizeof(unsigned long long)));
    PF_CREATE_CUDA(cudaHostAlloc((void**) &s->h_mail, 3 * sizeof(unsigned long long), cudaHostAllocMapped));
    memset(s->h_mail, 0, 3 * sizeof(unsigned long long));
    PF_CREATE_CUDA(cudaMallocHost((void**) &s->h_err, 2 * sizeof(unsigned long long)));
    PF_CREATE_CUDA(cudaMallocHost((void**) &s->h_cg, 2 * sizeof(CgState)));
    PF_CREATE_CUDA(cudaMallocHost((void**) &s->h_stats, CG_MAX_BLOCKS * sizeof(StatsPartial)));
    #undef PF_CREATE_CUDA
    *out = s;
    return PF_OK;
}

int pf_create(pf_solver** out, int device, int dtype, int nx, int ny)
{
    pf_slab full = {0, ny, 0, 0, 0};
    return pf_create_slab(out, device, dtype, nx, ny, &full);
}

void pf_destroy(pf_solver* s)
{
    if(!s) return;
    cudaSetDevice(s->device);
    if(s->stream) cudaStreamSynchronize(s->stream);
    for(int i = 0; i < PF_SCRATCH_SLOTS; i++) if(s->scratch[i]) cudaFree(s->scratch[i]);
    if(s->partials) cudaFree(s->partials);
    if(s->ll_slots) cudaFree(s->ll_slots);
    if(s->run_graph) cudaGraphExecDestroy(s->run_graph);
    if(s->run_progress) cudaFree(s->run_progress);
    if(s->halo_flags) cudaFree(s->halo_flags);
    if(s->d_cg) cudaFree(s->d_cg);
    if(s->d_stats) cudaFree(s->d_stats);
    if(s->d_err) cudaFree(s->d_err);
    if(s->staging) cudaFree(s->staging);
    if(s->h_err) cudaFreeHost(s->h_err);
    if(s->h_mail) cudaFreeHost(s->h_mail);
    if(s->h_cg) cudaFreeHost(s->h_cg);
    if(s->h_stats) cudaFreeHost(s->h_stats);
    if(s->aux_stream) { cudaStreamSynchronize(s->aux_stream); cudaStreamDestroy(s->aux_stream); }
    if(s->h2d_stream) { cudaStreamSynchronize(s->h2d_stream); cudaStreamDestroy(s->h2d_stream); }
    if(s->d2h_stream) { cudaStreamSynchronize(s->d2h_stream); cudaStreamDestroy(s->d2h_stream); }
    for(int i = 0; i < pf_solver::HOST_BANDS_MAX; i++) { if(s->ev_up[i]) cudaEventDestroy(s->ev_up[i]); if(s->ev_done[i]) cudaEventDestroy(s->ev_done[i]); }
    if(s->ev_host_begin) cudaEventDestroy(s->ev_host_begin);
    if(s->snap_stream) { cudaStreamSynchronize(s->snap_stream); cudaStreamDestroy(s->snap_stream); }
    for(auto& sl : s->snap)
    {
        if(sl.dev) cudaFree(sl.dev);
        if(sl.host) cudaFreeHost(sl.host);
        if(sl.staged) cudaEventDestroy(sl.staged);
        if(sl.landed) cudaEventDestroy(sl.landed);
    }
    if(s->d_rgba) cudaFree(s->d_rgba);
    if(s->stream) cudaStreamDestroy(s->stream);
    cudaGetLastError();
    delete s;
}

int pf_set_option(pf_solver* s, int option, int value)
{
    if(!s) return PF_ERR_INVALID;
    switch(option)
    {
        case PF_OPT_HOST_BANDS:    if(value < 1) return fail(s, PF_ERR_INVALID, "PF_OPT_HOST_BANDS: %d", value); s->host_bands = value; return PF_OK;
        case PF_OPT_SI_PER_KERNEL: s->force_per_kernel_cg = value != 0; return PF_OK;
        case PF_OPT_SI_GRID_SYNC:  s->plain_grid_sync = value != 0; return PF_OK;
        case PF_OPT_RUN_GRAPHS:    s->run_graphs_ok = value != 0; return PF_OK;
        case PF_OPT_SI_BLOCKS_PER_SM: if(value < 1 || value > 4) return fail(s, PF_ERR_INVALID, "PF_OPT_SI_BLOCKS_PER_SM: %d", value); s->si_blocks_per_sm = value; return PF_OK;
        case PF_OPT_HALO_IN_KERNEL: s->halo_in_kernel_ok = value != 0; return PF_OK;
        default: return fail(s, PF_ERR_INVALID, "pf_set_option: unknown option %d", option);
    }
}

int pf_reset_state(pf_solver* s)
{
    if(!s) return PF_ERR_INVALID;
    s->adaptive_tau = 0; s->last_num_steps = 0;
    return PF_OK;
}

int pf_field_alloc(pf_solver* s, void** dev_field)
{
    if(int rc = bind(s)) return rc;
    if(!dev_field) return fail(s, PF_ERR_INVALID, "pf_field_alloc: NULL out pointer");
    *dev_field = nullptr;
    PF_CUDA(s, cudaMalloc(dev_field, s->field_bytes()));
    PF_CUDA(s, cudaMemsetAsync(*dev_field, 0, s->field_bytes(), s->stream));      // REALLOC_ZERO, simulation.cu:92
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    return PF_OK;
}

int pf_field_free(pf_solver* s, void* dev_field)
{
    if(int rc = bind(s)) return rc;
    if(dev_field) { PF_CUDA(s, cudaStreamSynchronize(s->stream)); PF_CUDA(s, cudaFree(dev_field)); }
    return PF_OK;
}

int pf_field_rows(const pf_solver* s, int* rows_alloc, int* first_owned_row)
{
    if(!s) return PF_ERR_INVALID;
    if(rows_alloc) *rows_alloc = s->rows_alloc;
    if(first_owned_row) *first_owned_row = s->ghost_lo;
    return PF_OK;
}

static int ensure_staging(pf_solver* s, size_t bytes)
{
    if(s->staging_bytes < bytes)
    {
        if(s->staging) { PF_CUDA(s, cudaStreamSynchronize(s->stream)); PF_CUDA(s, cudaFree(s->staging)); s->staging = nullptr; s->staging_bytes = 0; }
        PF_CUDA(s, cudaMalloc(&s->staging, bytes));
        s->staging_bytes = bytes;
    }
    return PF_OK;
}

int pf_upload(pf_solver* s, void* dev, const void* host, size_t count, int host_dtype)
{
    if(int rc = bind(s)) return rc;
    if(!dev || !host) return fail(s, PF_ERR_INVALID, "pf_upload: NULL pointer");
    if(host_dtype != PF_F32 && host_dtype != PF_F64) return fail(s, PF_ERR_INVALID, "pf_upload: bad host dtype");
    if(host_dtype == s->dtype)
        PF_CUDA(s, cudaMemcpyAsync(dev, host, count * s->real_size(), cudaMemcpyHostToDevice, s->stream));
    else
    {
        // host -> staging -> field with the cast on the device, as sim_modify_T does (simulation.cu:1025-1031)
        size_t hs = pf_sizeof_real(host_dtype);
        if(int rc = ensure_staging(s, count * hs)) return rc;
        PF_CUDA(s, cudaMemcpyAsync(s->staging, host, count * hs, cudaMemcpyHostToDevice, s->stream));
        int blocks = elementwise_blocks(s, count, 256);
        if(s->dtype == PF_F64) k_convert<double, float><<<blocks, 256, 0, s->stream>>>(count, (const float*) s->staging, (double*) dev);
        else                   k_convert<float, double><<<blocks, 256, 0, s->stream>>>(count, (const double*) s->staging, (float*) dev);
        PF_LAUNCH_CHECK(s);
    }
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    return PF_OK;
}

int pf_download(pf_solver* s, const void* dev, void* host, size_t count, int host_dtype)
{
    if(int rc = bind(s)) return rc;
    if(!dev || !host) return fail(s, PF_ERR_INVALID, "pf_download: NULL pointer");
    if(host_dtype != PF_F32 && host_dtype != PF_F64) return fail(s, PF_ERR_INVALID, "pf_download: bad host dtype");
    if(host_dtype == s->dtype)
        PF_CUDA(s, cudaMemcpyAsync(host, dev, count * s->real_size(), cudaMemcpyDeviceToHost, s->stream));
    else
    {
        size_t hs = pf_sizeof_real(host_dtype);
        if(int rc = ensure_staging(s, count * hs)) return rc;
        int blocks = elementwise_blocks(s, count, 256);
        if(s->dtype == PF_F64) k_convert<float, double><<<blocks, 256, 0, s->stream>>>(count, (const double*) dev, (float*) s->staging);
        else                   k_convert<double, float><<<blocks, 256, 0, s->stream>>>(count, (const float*) dev, (double*) s->staging);
        PF_LAUNCH_CHECK(s);
        PF_CUDA(s, cudaMemcpyAsync(host, s->staging, count * hs, cudaMemcpyDeviceToHost, s->stream));
    }
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    return PF_OK;
}

int pf_upload_raw(pf_solver* s, void* dev, const void* host, size_t bytes)
{
    if(int rc = bind(s)) return rc;
    if(!dev || !host) return fail(s, PF_ERR_INVALID, "pf_upload_raw: NULL pointer");
    PF_CUDA(s, cudaMemcpyAsync(dev, host, bytes, cudaMemcpyHostToDevice, s->stream));
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    return PF_OK;
}

int pf_download_raw(pf_solver* s, const void* dev, void* host, size_t bytes)
{
    if(int rc = bind(s)) return rc;
    if(!dev || !host) return fail(s, PF_ERR_INVALID, "pf_download_raw: NULL pointer");
    PF_CUDA(s, cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, s->stream));
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    return PF_OK;
}

int pf_sync(pf_solver* s)
{
    if(int rc = bind(s)) return rc;
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    return PF_OK;
}

void* pf_stream(pf_solver* s) { return s ? (void*) s->stream : nullptr; }
void* pf_aux_stream(pf_solver* s) { return s ? (void*) s->aux_stream : nullptr; }

int64_t pf_kernel_launches(const pf_solver* s) { return s ? s->launches : 0; }

int pf_halo_region(const pf_solver* s, void* field, int side, void** send_ptr, void** recv_ptr, size_t* count)
{
    if(!s || !field) return PF_ERR_INVALID;
    const size_t rs = s->dtype == PF_F64 ? 8 : 4;
    const int g = s->slab.ghost;
    char* base = (char*) field;
    const size_t row_bytes = (size_t) s->nx * rs;
    if(count) *count = (size_t) g * (size_t) s->nx;
    if(side == 0)
    {
        if(!s->slab.has_lo) { if(send_ptr) *send_ptr = nullptr; if(recv_ptr) *recv_ptr = nullptr; if(count) *count = 0; return PF_OK; }
        if(recv_ptr) *recv_ptr = base;
        if(send_ptr) *send_ptr = base + row_bytes * (size_t) s->ghost_lo;
    }
    else
    {
        if(!s->slab.has_hi) { if(send_ptr) *send_ptr = nullptr; if(recv_ptr) *recv_ptr = nullptr; if(count) *count = 0; return PF_OK; }
        if(recv_ptr) *recv_ptr = base + row_bytes * (size_t)(s->ghost_lo + s->slab.ny_local);
        if(send_ptr) *send_ptr = base + row_bytes * (size_t)(s->ghost_lo + s->slab.ny_local - g);
    }
    return PF_OK;
}

} // extern "C"

// -------------------------------------------------------------------------------------------------------
// Fused explicit step launchers
// -------------------------------------------------------------------------------------------------------
// ---- streaming kernel (pf_stream.cuh): launch geometry -------------------------------------------------
// CTA = 4 warps (128 columns, 128 - 2*NST of them interior). Resident CTAs per SM by register budget:
// Euler 8 (<= 64 registers), the multi-stage schemes 4 (<= 128 registers).
enum { STREAM_WARPS = 4 };
// Resident CTAs per SM the streaming kernels are compiled for. Merson in fp64 needs 148 - 164 registers: at 4 CTAs (128 registers)
// it spills 64 - 136 bytes per thread into the row loop; at 3 CTAs it does not (2048^2 adaptive step 129.7 -> 121.5 us, same bits).
template <typename Real, int SCHEME, bool ANISO> struct StreamOcc {
    static constexpr int MINB = SCHEME == SCHEME_EULER ? (ANISO ? 6 : 8) : (SCHEME == SCHEME_RKM && sizeof(Real) == 8 ? 3 : 4);
};

// Number of row chunks: every CTA marches down its chunk plus a pipeline fill of 2(NST-1) rows, so few long
// chunks waste the least work, but the grid should fill the machine's resident-CTA slots in whole waves. Cost model
// (fitted to kbench runs on B200): a wave costs iterations x (a + b x resident CTAs per SM).
static int stream_chunks(int sm_count, int minb, int strips, int rows, int fill)
{
    const long long slots = (long long) sm_count * minb;
    double best = 1e300; int best_n = 1;
    const int max_n = std::max(1, rows / 4);
    for(int n = 1; n <= max_n; n++)
    {
        const int chunk = (rows + n - 1) / n;                         // the longest chunk sets the time
        const int nch = (rows + chunk - 1) / chunk;                   // the fewest chunks that are no longer than that
        const long long ctas = (long long) strips * nch;
        const long long full = ctas / slots, rest = ctas % slots;
        const double its = chunk + fill + 3.0;                       // the checked first/last groups run slower
        const double occ_rest = std::min<double>(minb, std::ceil((double) rest / sm_count));
        const double cost = its * (full * (0.6 + 0.145 * minb) + (rest ? 0.6 + 0.145 * occ_rest : 0.0));
        if(cost < best) { best = cost; best_n = nch; }
    }
    return best_n;
}

// ANISO: bit 0 = anisotropy, bit 1 = the corrector-guess form of the right-hand side (simulation.cu:221-225; single-step launches only)
template <typename Real, int SCHEME, int ANISO, int MINB = StreamOcc<Real, SCHEME, (ANISO & 1) != 0>::MINB>
static int launch_stream_inst(pf_solver* s, FusedArgs<Real> a, int inst_bit)
{
    constexpr int NST = SchemeStages<SCHEME>::value, W = STREAM_WARPS;
    constexpr int EWI = 32 * W - 2 * NST;
    static_assert(EWI > 0, "strip too narrow for the stage count");
    auto kern = k_stream_step<Real, SCHEME, ANISO, W, MINB>;
    constexpr bool HAS_RUN = SCHEME == SCHEME_RK4 && ANISO < 2;             // the only scheme with a multi-step kernel (see below)
    const size_t smem = stream_smem_bytes<Real, SCHEME, W>();
    if(!(s->stream_attr_done & (1u << inst_bit)))
    {
        PF_CUDA(s, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        PF_CUDA(s, cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        if constexpr (HAS_RUN)
        {
            auto kern_run = k_stream_run<Real, SCHEME, ANISO, W, MINB>;
            PF_CUDA(s, cudaFuncSetAttribute(kern_run, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
            PF_CUDA(s, cudaFuncSetAttribute(kern_run, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        }
        s->stream_attr_done |= 1u << inst_bit;
    }
    const int own_rows = a.own_y1 - a.own_y0;
    s->run_used = false;
    if(own_rows <= 0) return PF_OK;
    a.strips = stream_strips(a.nx, EWI, NST, a.bc == PF_BOUNDARY_PERIODIC);
    a.nch = stream_chunks(s->sm_count, MINB, a.strips, own_rows, StreamGeom<SCHEME>::FILL);
    const long long units = (long long) a.strips * a.nch;
    cudaStream_t stream = s->launch_stream ? s->launch_stream : s->stream;
    // Multi-step launch (pf_run). Measured on B200 against single-step launches (tools/kbench.cu run_multi, fp64): explicit-rk4
    // 2048^2 -1 %, 1024^2 -2 %, 512^2 -10 %, 256^2 -6 %; the one-stage Euler step gains nothing from it (2048^2 +1 %, 512^2 +28 %:
    // the neighbour hand-shake costs what the launch gap costs), so it keeps its single-step launches.
    if constexpr (HAS_RUN) if(s->run_steps > 1)
    {
        // every unit must be resident at once and a tile may only depend on its 8 neighbours
        auto kern_run = k_stream_run<Real, SCHEME, ANISO, W, MINB>;
        if(!s->stream_resident[inst_bit])
        {
            int coop = 0, per_sm = 0;
            PF_CUDA(s, cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, s->device));
            if(coop) PF_CUDA(s, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern_run, 32 * W, smem));
            s->stream_resident[inst_bit] = coop && per_sm > 0 ? per_sm : -1;
        }
        const bool whole_grid = s->is_full_grid() && s->win_y1 <= s->win_y0;
        // one wave, and chunks of at least 2 x NST rows (a tile then only depends on its 8 neighbours)
        const long long slots = s->stream_resident[inst_bit] > 0 ? (long long) s->stream_resident[inst_bit] * s->sm_count : 0;
        if(whole_grid && units <= slots && own_rows / a.nch >= 2 * NST)
        {
            if(s->run_progress_units < units)
            {
                if(s->run_progress) { PF_CUDA(s, cudaStreamSynchronize(s->stream)); PF_CUDA(s, cudaFree(s->run_progress)); s->run_progress = nullptr; }
                PF_CUDA(s, cudaMalloc((void**) &s->run_progress, (size_t)(units + 1) * sizeof(unsigned)));
                s->run_progress_units = (int) units;
                s->run_base = 0;
            }
            if(s->run_base == 0 || s->run_base > 0x7f000000u)             // first use / before the counter can wrap: start over from zeroed words
            {
                PF_CUDA(s, cudaMemsetAsync(s->run_progress, 0, (size_t)(s->run_progress_units + 1) * sizeof(unsigned), stream));
                s->run_base = 1;
            }
            StreamRun r; r.n_steps = s->run_steps; r.base = s->run_base; r.progress = s->run_progress; r.units = s->run_progress_units;
            s->run_base += (unsigned) s->run_steps;
            void* args[] = {&a, &r};
            PF_CUDA(s, cudaLaunchCooperativeKernel((const void*) kern_run, dim3((unsigned) units), dim3(32 * W), args, smem, stream));
            PF_LAUNCH_CHECK(s);
            s->run_used = true;
            return PF_OK;
        }
    }
    if constexpr (SCHEME != SCHEME_EULER && ANISO < 2)
    {
        // a linked slab (pf_halo_link), whole-slab launch: the halo hand-over happens inside the kernel
        if(a.halo_flags && s->halo_in_kernel_ok && own_rows / a.nch >= a.halo_g)
        {
            auto kern_halo = k_stream_step_halo<Real, SCHEME, ANISO, W, MINB>;
            if(!(s->halo_attr_done & (1u << inst_bit)))
            {
                PF_CUDA(s, cudaFuncSetAttribute(kern_halo, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
                PF_CUDA(s, cudaFuncSetAttribute(kern_halo, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
                s->halo_attr_done |= 1u << inst_bit;
            }
            kern_halo<<<(unsigned) units, 32 * W, smem, stream>>>(a);
            PF_LAUNCH_CHECK(s);
            s->halo_in_kernel = true;
            return PF_OK;
        }
    }
    if(a.halo_flags) if(int rc = halo_wait_both(s)) return rc;           // linked slab, no in-kernel hand-over: the caller pushes afterwards
    kern<<<(unsigned) units, 32 * W, smem, stream>>>(a);
    PF_LAUNCH_CHECK(s);
    return PF_OK;
}

// ---- Euler tile kernel (pf_euler.cuh): launch geometry, measured with tools/kbench.cu on B200 at 2048^2 ---------------------------
// fp64: R = 8 rows per tile, 7 CTAs of 128 threads per SM (72 registers): 26.8 us (R = 8 at 6 or 8 CTAs, R = 6, 10, 12, 16, 96- / 192- /
// 256-thread tiles: 27.8-31.4); fp32: 8 CTAs: 19.0 us; with anisotropy (atan2f / cosf per cell): R = 4, 6 CTAs: 41.0 us.
// GENERAL: corrector guess / source term / corrector iteration / different boundary types of Phi and T (smaller tiles: more state).
template <typename Real, bool ANISO, bool GENERAL = false>
static int launch_euler_inst(pf_solver* s, const FusedArgs<Real>& a, int inst_bit)
{
    constexpr int EW = 128, R = (ANISO || GENERAL) ? 4 : 8, MINB = (ANISO || GENERAL) ? 6 : (sizeof(Real) == 8 ? 7 : 8);
    auto kern = k_euler_tile<Real, ANISO, GENERAL, EW, R, MINB>;
    const size_t smem = euler_tile_smem_bytes<Real, EW, R>();
    if(!(s->stream_attr_done & (1u << inst_bit)))
    {
        PF_CUDA(s, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        s->stream_attr_done |= 1u << inst_bit;
    }
    const int own_rows = a.own_y1 - a.own_y0;
    s->run_used = false;
    if(own_rows <= 0) return PF_OK;
    cudaStream_t stream = s->launch_stream ? s->launch_stream : s->stream;
    dim3 grid((unsigned)((a.nx + (EW - 2 * EULER_HALO) - 1) / (EW - 2 * EULER_HALO)), (unsigned)((own_rows + R - 1) / R));
    if(a.halo_flags) if(int rc = halo_wait_both(s)) return rc;           // linked slab: the caller pushes afterwards
    kern<<<grid, EW, smem, stream>>>(a);
    PF_LAUNCH_CHECK(s);
    return PF_OK;
}

// One launch = one whole time step (or one Merson attempt): the Euler tile kernel (pf_euler.cuh) for the one-stage scheme, the
// streaming kernel (pf_stream.cuh) for RK4 and Merson, with instantiations of their own for the corrector-guess form of the
// right-hand side (simulation.cu:221-225). (Euler with the guess: the GENERAL tile kernel, euler_step_based.)
template <typename Real>
static int launch_fused(pf_solver* s, int scheme, bool aniso, const FusedArgs<Real>& a)
{
    if(a.k.do_corrector_guess && scheme != SCHEME_EULER)
    {
        const int bit = 17 + (sizeof(Real) == 8 ? 0 : 4) + (scheme == SCHEME_RKM ? 2 : 0) + (aniso ? 1 : 0);
        switch((scheme == SCHEME_RKM ? 2 : 0) + (aniso ? 1 : 0))
        {
            case 0: return launch_stream_inst<Real, SCHEME_RK4, 2>(s, a, bit);
            case 1: return launch_stream_inst<Real, SCHEME_RK4, 3>(s, a, bit);
            case 2: return launch_stream_inst<Real, SCHEME_RKM, 2>(s, a, bit);
            default: return launch_stream_inst<Real, SCHEME_RKM, 3>(s, a, bit);
        }
    }
    const int base = (sizeof(Real) == 8 ? 0 : 6) + scheme * 2 + (aniso ? 1 : 0);
    switch(scheme * 2 + (aniso ? 1 : 0))
    {
        case 0: return launch_euler_inst<Real, false>(s, a, base);
        case 1: return launch_euler_inst<Real, true>(s, a, base);
        case 2:
            // explicit-rk4 without anisotropy, fp64 (the benchmark family): compiled for 5 CTAs per SM as well (96 registers, 16 bytes
            // of spills) - taken when the grid is ONE wave at 5 per SM (2048^2: 72.6 against 75.1 us; 6 per SM: 76.0, 3: 85.8);
            // the many-wave 16384 x 2048 slab stays at 4 (0.472 against 0.476 ms), and so do grids whose chunks would get shorter than 40
            // rows at 5 per SM (1024^2: 28.0 us at 4, 30.9 at 5; 512^2: 24.5 / 30.1 - the pipeline fill of 9 rows per chunk)
            if constexpr (sizeof(Real) == 8)
            {
                constexpr int NST = SchemeStages<SCHEME_RK4>::value, EWI = 32 * STREAM_WARPS - 2 * NST;
                const int rows = a.own_y1 - a.own_y0;
                if(rows > 0)
                {
                    const int strips = stream_strips(a.nx, EWI, NST, a.bc == PF_BOUNDARY_PERIODIC);
                    const int nch = stream_chunks(s->sm_count, 5, strips, rows, StreamGeom<SCHEME_RK4>::FILL);
                    if((long long) strips * nch <= (long long) s->sm_count * 5 && rows / nch >= 40) return launch_stream_inst<Real, SCHEME_RK4, false, 5>(s, a, 16);
                }
            }
            return launch_stream_inst<Real, SCHEME_RK4, false>(s, a, base);
        case 3: return launch_stream_inst<Real, SCHEME_RK4, true>(s, a, base);
        case 4: return launch_stream_inst<Real, SCHEME_RKM, false>(s, a, base);
        default: return launch_stream_inst<Real, SCHEME_RKM, true>(s, a, base);
    }
}

template <typename Real>
static FusedArgs<Real> make_fused_args(const pf_solver* s, const pf_params& p, const void* F, const void* U, void* nF, void* nU)
{
    FusedArgs<Real> a;
    memset(&a, 0, sizeof a);
    a.F = (const Real*) F; a.U = (const Real*) U; a.out_F = (Real*) nF; a.out_U = (Real*) nU;
    a.nx = s->nx; a.ny_global = s->ny;
    a.row0 = s->slab.y0 - s->ghost_lo; a.rows_local = s->rows_alloc;
    a.own_y0 = s->slab.y0; a.own_y1 = s->slab.y0 + s->slab.ny_local;
    if(s->win_y1 > s->win_y0) { a.own_y0 = std::max(a.own_y0, s->win_y0); a.own_y1 = std::min(a.own_y1, s->win_y1); }
    a.bc = p.Phi_boundary; a.bc_T = p.T_boundary; a.U_base = nullptr;
    a.wrap_y = (p.Phi_boundary == PF_BOUNDARY_PERIODIC && !s->slab.has_lo && !s->slab.has_hi) ? 1 : 0;
    a.k = make_consts<Real>(p);
    a.err_bits = s->d_err;
    return a;
}

// ---- linked slabs (pf_halo_link): every whole-slab step of the fixed-step explicit solvers exchanges its edge rows ------------------
static bool halo_active(const pf_solver* s) { return s->halo.on && s->win_y1 <= s->win_y0 && !s->launch_stream; }

static int halo_index(const pf_solver* s, const void* field)
{
    for(int i = 0; i < s->halo.n_fields; i++) if(s->halo.local[i] == field) return i;
    return -1;
}

// Fills the in-kernel hand-over of FusedArgs for the output fields nF / nU; false if they are not linked fields.
template <typename Real>
static bool halo_fill(pf_solver* s, FusedArgs<Real>& a, void* nF, void* nU)
{
    const int iF = halo_index(s, nF), iU = halo_index(s, nU);
    if(iF < 0 || iU < 0) return false;
    a.halo_e = s->halo.e; a.halo_g = s->slab.ghost; a.halo_flags = s->halo_flags;
    for(int side = 0; side < 2; side++)
    {
        const bool there = s->halo.peer_flags[side] != nullptr;
        const size_t off = s->halo.first_row[side] * (size_t) s->nx;
        a.peer_F[side] = there ? (Real*) s->halo.peer[side][iF] + off : nullptr;
        a.peer_U[side] = there ? (Real*) s->halo.peer[side][iU] + off : nullptr;
        a.peer_flag[side] = there ? s->halo.peer_flags[side] + (side == 0 ? HALO_FLAG_FROM_HI : HALO_FLAG_FROM_LO) : nullptr;
    }
    return true;
}

// The same exchange with separate launches around a step kernel that cannot do it itself (Euler tile kernel, chunks shorter than the
// ghost depth): wait kernels before ...
static int halo_wait_both(pf_solver* s)
{
    for(int side = 0; side < 2; side++)
        if(s->halo.peer_flags[side]) if(int rc = pf_halo_wait_async(s, side, s->halo.e, 0)) return rc;
    return PF_OK;
}
// ... and push kernels after.
static int halo_push_both(pf_solver* s, void* nF, void* nU)
{
    const int iF = halo_index(s, nF), iU = halo_index(s, nU);
    for(int side = 0; side < 2; side++)
    {
        if(!s->halo.peer_flags[side]) continue;
        const void* mine[2] = {nF, nU};
        void* theirs[2] = {s->halo.peer[side][iF], s->halo.peer[side][iU]};
        if(int rc = pf_halo_push_async(s, side, 2, mine, theirs, s->halo.first_row[side], s->halo.peer_flags[side],
                                       side == 0 ? HALO_FLAG_FROM_HI : HALO_FLAG_FROM_LO, s->halo.e + 1, 0)) return rc;
    }
    return PF_OK;
}

static bool fused_eligible(const pf_solver* s, const pf_params& p)
{
    // the single-kernel paths share one load geometry for both fields: Phi and T both periodic or both not (any Dirichlet / Neumann mix)
    const bool perF = p.Phi_boundary == PF_BOUNDARY_PERIODIC, perT = p.T_boundary == PF_BOUNDARY_PERIODIC;
    if(perF != perT) return false;
    if(p.do_exact) return false;                                      // nonzero source term fu in every stage: general per-stage path
    if(perF && (s->slab.has_lo != s->slab.has_hi)) return false;
    return true;
}

// -------------------------------------------------------------------------------------------------------
// General per-stage path (any boundary mix; single GPU full grid only)
// -------------------------------------------------------------------------------------------------------
template <typename Real>
static int launch_rhs_general(pf_solver* s, const pf_params& p, bool is_euler, int n_states, const Real* w,
                              const Real* const* Fs, const Real* const* Us, Real* out_F, Real* out_U)
{
    BlendArgs<Real> b; memset(&b, 0, sizeof b);
    b.n_states = n_states;
    for(int i = 0; i < n_states; i++) { b.w[i] = w[i]; b.F[i] = Fs[i]; b.U[i] = Us[i]; }
    KConsts<Real> k = make_consts<Real>(p);
    dim3 block(32, 8), grid((unsigned)((s->nx + 31) / 32), (unsigned)((s->ny + 7) / 8));
    if(p.S != 0) k_rhs_general<Real, true><<<grid, block, 0, s->stream>>>(b, k, s->nx, s->ny, p.Phi_boundary, p.T_boundary, is_euler ? 1 : 0, out_F, out_U);
    else         k_rhs_general<Real, false><<<grid, block, 0, s->stream>>>(b, k, s->nx, s->ny, p.Phi_boundary, p.T_boundary, is_euler ? 1 : 0, out_F, out_U);
    PF_LAUNCH_CHECK(s);
    return PF_OK;
}

// The Euler tile kernel serves every parameter combination whose loads can share one geometry: Phi and T both periodic or both not.
static bool euler_tile_eligible(const pf_solver* s, const pf_params& p)
{
    const bool perF = p.Phi_boundary == PF_BOUNDARY_PERIODIC, perT = p.T_boundary == PF_BOUNDARY_PERIODIC;
    if(perF != perT) return false;
    if(perF && (s->slab.has_lo != s->slab.has_hi)) return false;
    if(p.do_exact && !s->is_full_grid()) return false;
    return true;
}

// explicit_solver_euler_step_based: reference simulation.cu:283-311. One launch in every case the tile kernel is eligible for:
// the plain step (lean RHS), and - GENERAL instantiation - the corrector-guess RHS, the source term of do_exact, different
// non-periodic boundary types for Phi and T, and the corrector iteration y(n+1) = {F, U_base} + dt f(F, U).
template <typename Real>
static int euler_step_based(pf_solver* s, const pf_params& p, const Real* F, const Real* U, const Real* U_base, Real* nF, Real* nU, pf_step_info* info)
{
    if(euler_tile_eligible(s, p))
    {
        FusedArgs<Real> a = make_fused_args<Real>(s, p, F, U, nF, nU);
        const bool general = U != U_base || p.do_corrector_guess || p.do_exact || p.Phi_boundary != p.T_boundary;
        if(U != U_base) a.U_base = U_base;
        const int bit = 12 + (sizeof(Real) == 8 ? 0 : 2) + (p.S != 0 ? 1 : 0);
        const bool hx = U == U_base && halo_active(s) && halo_fill<Real>(s, a, nF, nU);
        int rc;
        if(!general)     rc = launch_fused<Real>(s, SCHEME_EULER, p.S != 0, a);
        else if(p.S != 0) rc = launch_euler_inst<Real, true, true>(s, a, bit);
        else              rc = launch_euler_inst<Real, false, true>(s, a, bit);
        if(rc) return rc;
        if(hx)
        {
            if(int rc2 = halo_push_both(s, nF, nU)) return rc2;
            s->halo.e++;
        }
    }
    else if(U == U_base)
    {
        if(!s->is_full_grid()) return fail(s, PF_ERR_UNSUPPORTED, "periodic mixed with non-periodic boundary types is single-GPU only");
        Real one = 1; const Real* Fs[1] = {F}; const Real* Us[1] = {U};
        if(int rc = launch_rhs_general<Real>(s, p, true, 1, &one, Fs, Us, nF, nU)) return rc;
    }
    else
    {
        if(!s->is_full_grid()) return fail(s, PF_ERR_UNSUPPORTED, "periodic mixed with non-periodic boundary types is single-GPU only");
        void *dF, *dU;
        if(int rc = get_scratch(s, 0, &dF)) return rc;
        if(int rc = get_scratch(s, 1, &dU)) return rc;
        Real one = 1; const Real* Fs[1] = {F}; const Real* Us[1] = {U};
        if(int rc = launch_rhs_general<Real>(s, p, false, 1, &one, Fs, Us, (Real*) dF, (Real*) dU)) return rc;
        size_t N = s->field_elems();
        k_euler_corrector_axpy<Real><<<elementwise_blocks(s, N, 256), 256, 0, s->stream>>>(N, (Real) p.dt, F, U_base, (const Real*) dF, (const Real*) dU, nF, nU);
        PF_LAUNCH_CHECK(s);
    }
    if(info) { info->Phi_iters = 1; info->T_iters = 1; }
    return PF_OK;
}

// explicit_solver_rk4_step: reference simulation.cu:313-348.
template <typename Real>
static int rk4_step(pf_solver* s, const pf_params& p, const Real* F, const Real* U, Real* nF, Real* nU, pf_step_info* info)
{
    const Real dt = (Real) p.dt;
    if(fused_eligible(s, p))
    {
        FusedArgs<Real> a = make_fused_args<Real>(s, p, F, U, nF, nU);
        a.w[0] = dt / 2; a.w[1] = dt / 2; a.w[2] = dt; a.w[3] = dt / 6;
        const bool hx = halo_active(s) && halo_fill<Real>(s, a, nF, nU);
        s->halo_in_kernel = false;
        if(int rc = launch_fused<Real>(s, SCHEME_RK4, p.S != 0, a)) return rc;
        if(hx)
        {
            if(!s->halo_in_kernel) if(int rc = halo_push_both(s, nF, nU)) return rc;
            s->halo.e++;
        }
    }
    else
    {
        if(!s->is_full_grid()) return fail(s, PF_ERR_UNSUPPORTED, "mixed boundary types are single-GPU only");
        Real* k[8];
        for(int i = 0; i < 8; i++) { void* q; if(int rc = get_scratch(s, i, &q)) return rc; k[i] = (Real*) q; }
        Real *k1F = k[0], *k2F = k[1], *k3F = k[2], *k4F = k[3], *k1U = k[4], *k2U = k[5], *k3U = k[6], *k4U = k[7];
        { Real w[1] = {1}; const Real* Fs[1] = {F}; const Real* Us[1] = {U};
          if(int rc = launch_rhs_general<Real>(s, p, false, 1, w, Fs, Us, k1F, k1U)) return rc; }
        { Real w[2] = {1, dt / 2}; const Real* Fs[2] = {F, k1F}; const Real* Us[2] = {U, k1U};
          if(int rc = launch_rhs_general<Real>(s, p, false, 2, w, Fs, Us, k2F, k2U)) return rc; }
        { Real w[2] = {1, dt / 2}; const Real* Fs[2] = {F, k2F}; const Real* Us[2] = {U, k2U};
          if(int rc = launch_rhs_general<Real>(s, p, false, 2, w, Fs, Us, k3F, k3U)) return rc; }
        { Real w[2] = {1, dt}; const Real* Fs[2] = {F, k3F}; const Real* Us[2] = {U, k3U};
          if(int rc = launch_rhs_general<Real>(s, p, false, 2, w, Fs, Us, k4F, k4U)) return rc; }
        size_t N = s->field_elems();
        int blocks = elementwise_blocks(s, N, 256);
        k_rk4_combine<Real><<<blocks, 256, 0, s->stream>>>(N, dt / 6, F, k1F, k2F, k3F, k4F, nF); PF_LAUNCH_CHECK(s);
        k_rk4_combine<Real><<<blocks, 256, 0, s->stream>>>(N, dt / 6, U, k1U, k2U, k3U, k4U, nU); PF_LAUNCH_CHECK(s);
    }
    if(info) { info->Phi_iters = 1; info->T_iters = 1; }
    return PF_OK;
}

// Decodes the atomicMax'ed bit patterns of the two error norms.
template <typename Real> static Real decode_err(unsigned long long bits);
template <> double decode_err<double>(unsigned long long bits) { double d; memcpy(&d, &bits, 8); return d; }
template <> float decode_err<float>(unsigned long long bits) { uint32_t u = (uint32_t) bits; float f; memcpy(&f, &u, 4); return f; }

static inline void pf_cpu_relax()
{
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#endif
}

static inline double pf_pow(double a, double b) { return pow(a, b); }
static inline float pf_pow(float a, float b) { return powf(a, b); }

// explicit_solver_rk4_adaptive_step (Runge-Kutta-Merson): reference simulation.cu:350-497, split into the three
// phases a multi-rank caller needs (begin / attempt / decide); pf_step runs them back to back.
// The step-size controller runs on the host in Real arithmetic with the same libm pow as the reference's host code.

// :363-386 - the carried step size and, on the per-stage path, k1 (computed once per step, outside the attempt loop).
template <typename Real>
static int rkm_begin(pf_solver* s, const pf_params& p, const Real* F, const Real* U, double* tau_out)
{
    if(s->adaptive_tau <= 0) s->adaptive_tau = (double)(Real) p.dt;        // :363-365
    *tau_out = s->adaptive_tau;
    s->rkm_fused = fused_eligible(s, p);
    if(!s->rkm_fused)
    {
        if(!s->is_full_grid()) return fail(s, PF_ERR_UNSUPPORTED, "mixed boundary types are single-GPU only");
        Real* k[10];
        for(int i = 0; i < 10; i++) { void* q; if(int rc = get_scratch(s, i, &q)) return rc; k[i] = (Real*) q; }
        Real w[1] = {1}; const Real* Fs[1] = {F}; const Real* Us[1] = {U};
        if(int rc = launch_rhs_general<Real>(s, p, false, 1, w, Fs, Us, k[0], k[5])) return rc;      // k1, :386
    }
    return PF_OK;
}

// :406-435 - k2..k5 for the step size tau and the two max-norms, left on the device in s->d_err (as bit patterns).
// On the fused path the kernel also writes the candidate result y + tau/6 (k1 + 4 k4 + k5) into nF/nU (:471-476).
template <typename Real>
static int rkm_attempt(pf_solver* s, const pf_params& p, double tau_d, const Real* F, const Real* U, Real* nF, Real* nU)
{
    const Real tau = (Real) tau_d;
    const bool handover = s->rkm_handover && s->rkm_fused && !s->rkm_keep_errors && s->is_full_grid() && s->win_y1 <= s->win_y0;
    if(handover ? s->d_err_dirty : !s->rkm_keep_errors) PF_CUDA(s, cudaMemsetAsync(s->d_err, 0, 3 * sizeof(unsigned long long), s->stream));
    s->d_err_dirty = !handover;
    if(s->rkm_fused)
    {
        FusedArgs<Real> a = make_fused_args<Real>(s, p, F, U, nF, nU);
        if(handover)
        {
            unsigned long long* dev_mail = nullptr;
            PF_CUDA(s, cudaHostGetDevicePointer((void**) &dev_mail, s->h_mail, 0));
            a.err_host = dev_mail; a.err_ticket = (unsigned*)(s->d_err + 2); a.err_seq = ++s->mail_seq;
        }
        a.w[0] = tau / 3; a.w[1] = tau / 6; a.w[2] = tau / 8; a.w[3] = tau * 3 / 8;
        a.w[4] = tau / 2; a.w[5] = -tau * 3 / 2; a.w[6] = tau * 2; a.w[7] = tau / 6;
        // linked slab: every attempt hands its candidate rows to the neighbours (a rejected attempt's rows are overwritten by the next one's)
        const bool hx = halo_active(s) && halo_fill<Real>(s, a, nF, nU);
        s->halo_in_kernel = false;
        if(int rc = launch_fused<Real>(s, SCHEME_RKM, p.S != 0, a)) return rc;
        if(hx)
        {
            if(!s->halo_in_kernel) if(int rc = halo_push_both(s, nF, nU)) return rc;
            s->halo.e++;
        }
        return PF_OK;
    }
    Real* k[10];
    for(int i = 0; i < 10; i++) k[i] = (Real*) s->scratch[i];
    Real *k1F = k[0], *k2F = k[1], *k3F = k[2], *k4F = k[3], *k5F = k[4], *k1U = k[5], *k2U = k[6], *k3U = k[7], *k4U = k[8], *k5U = k[9];
    { Real w[2] = {1, tau / 3}; const Real* Fs[2] = {F, k1F}; const Real* Us[2] = {U, k1U};
      if(int rc = launch_rhs_general<Real>(s, p, false, 2, w, Fs, Us, k2F, k2U)) return rc; }
    { Real w[3] = {1, tau / 6, tau / 6}; const Real* Fs[3] = {F, k1F, k2F}; const Real* Us[3] = {U, k1U, k2U};
      if(int rc = launch_rhs_general<Real>(s, p, false, 3, w, Fs, Us, k3F, k3U)) return rc; }
    { Real w[3] = {1, tau / 8, tau * 3 / 8}; const Real* Fs[3] = {F, k1F, k3F}; const Real* Us[3] = {U, k1U, k3U};
      if(int rc = launch_rhs_general<Real>(s, p, false, 3, w, Fs, Us, k4F, k4U)) return rc; }
    { Real w[4] = {1, tau / 2, -tau * 3 / 2, tau * 2}; const Real* Fs[4] = {F, k1F, k3F, k4F}; const Real* Us[4] = {U, k1U, k3U, k4U};
      if(int rc = launch_rhs_general<Real>(s, p, false, 4, w, Fs, Us, k5F, k5U)) return rc; }
    const size_t N = s->field_elems();
    const int blocks = elementwise_blocks(s, N, 256);
    k_rkm_error<Real><<<blocks, 256, 0, s->stream>>>(N, k1F, k3F, k4F, k5F, s->d_err + 0); PF_LAUNCH_CHECK(s);
    k_rkm_error<Real><<<blocks, 256, 0, s->stream>>>(N, k1U, k3U, k4U, k5U, s->d_err + 1); PF_LAUNCH_CHECK(s);
    return PF_OK;
}

// :437-467 - the accept test and the step-size controller, from the (globally reduced) max-norm bit patterns.
// *stop: 0 = try again with *next_tau, 1 = accepted, 2 = the min_dt floor was reached (the reference `break`s there,
// accepted or not, without counting the attempt).
template <typename Real>
static void rkm_decide(const pf_params& p, double tau_d, unsigned long long bits_F, unsigned long long bits_U,
                       int* converged, int* stop, double* next_tau, double* eps_F, double* eps_U)
{
    const Real tau = (Real) tau_d;
    const Real max_F = decode_err<Real>(bits_F);
    const Real max_U = decode_err<Real>(bits_U);
    const Real epsilon_F = tau / 3 * max_F;                                 // :437-438
    const Real epsilon_U = tau / 3 * max_U;
    *converged = ((double) epsilon_F < p.Phi_tolerance && (double) epsilon_U < p.T_tolerance) ? 1 : 0;   // :456
    const double e_max = (double)(epsilon_F > epsilon_U ? epsilon_F : epsilon_U);
    const Real epsilon = (Real)(e_max > 1e-20 ? e_max : 1e-20);             // :459
    const double d_min = p.Phi_tolerance < p.T_tolerance ? p.Phi_tolerance : p.T_tolerance;
    const Real delta = (Real)(d_min > 1e-20 ? d_min : 1e-20);               // :460
    Real nt = pf_pow(delta / epsilon, (Real) 0.2) * 4 / 5 * tau;            // :462
    nt = (Real)((double) nt > p.min_dt ? (double) nt : p.min_dt);           // :463
    const bool floor_hit = (double) nt <= p.min_dt && (double) tau <= p.min_dt;   // :466
    *stop = floor_hit ? 2 : (*converged ? 1 : 0);
    *next_tau = (double) nt;
    *eps_F = (double) epsilon_F; *eps_U = (double) epsilon_U;
}

// :471-476 on the per-stage path (the fused kernel has already written the result).
template <typename Real>
static int rkm_finish(pf_solver* s, double used_tau_d, const Real* F, const Real* U, Real* nF, Real* nU)
{
    if(s->rkm_fused) return PF_OK;
    const Real used_tau = (Real) used_tau_d;
    Real* k[10];
    for(int i = 0; i < 10; i++) k[i] = (Real*) s->scratch[i];
    const size_t N = s->field_elems();
    const int blocks = elementwise_blocks(s, N, 256);
    k_rkm_combine<Real><<<blocks, 256, 0, s->stream>>>(N, used_tau / 6, F, k[0], k[3], k[4], nF); PF_LAUNCH_CHECK(s);
    k_rkm_combine<Real><<<blocks, 256, 0, s->stream>>>(N, used_tau / 6, U, k[5], k[8], k[9], nU); PF_LAUNCH_CHECK(s);
    return PF_OK;
}

template <typename Real>
static int rk4_adaptive_step(pf_solver* s, const pf_params& p, const Real* F, const Real* U, Real* nF, Real* nU, pf_step_info* info, double* dt_out)
{
    double tau = 0;
    if(int rc = rkm_begin<Real>(s, p, F, U, &tau)) return rc;
    const int max_iters = std::max(std::max(p.T_max_iters, p.Phi_max_iters), 1);  // :390
    int converged = 0, stop = 0, i = 0;
    double used_tau = tau, eps_F = 0, eps_U = 0;
    for(; i < max_iters && !converged; i++)
    {
        s->rkm_handover = true;
        const unsigned long long seq_before = s->mail_seq;
        int rc = rkm_attempt<Real>(s, p, tau, F, U, nF, nU);
        s->rkm_handover = false;
        if(rc) return rc;
        unsigned long long bits_F, bits_U;
        if(s->mail_seq != seq_before)
        {
            // the kernel's last unit writes the norms into mapped host memory: spin on the sequence number (1-2 us after the
            // kernel's end instead of ~10 for a copy + stream synchronisation)
            volatile unsigned long long* m = s->h_mail;
            for(unsigned long spins = 1; m[2] != s->mail_seq; spins++)
            {
                pf_cpu_relax();
                if((spins & 0x3ffff) == 0)
                {
                    const cudaError_t q = cudaStreamQuery(s->stream);
                    if(q == cudaSuccess && m[2] != s->mail_seq) { __sync_synchronize(); if(m[2] == s->mail_seq) break; return fail(s, PF_ERR_CUDA, "adaptive step: the attempt finished without reporting its error norms"); }
                    if(q != cudaSuccess && q != cudaErrorNotReady) return fail(s, PF_ERR_CUDA, "adaptive step: %s", cudaGetErrorString(q));
                }
            }
            __sync_synchronize();
            bits_F = m[0]; bits_U = m[1];
        }
        else
        {
            PF_CUDA(s, cudaMemcpyAsync(s->h_err, s->d_err, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s->stream));
            PF_CUDA(s, cudaStreamSynchronize(s->stream));
            bits_F = s->h_err[0]; bits_U = s->h_err[1];
        }
        used_tau = tau;
        rkm_decide<Real>(p, used_tau, bits_F, bits_U, &converged, &stop, &tau, &eps_F, &eps_U);
        if(stop == 2) break;                                                // :466: `break` skips the loop's i++
    }
    if(int rc = rkm_finish<Real>(s, used_tau, F, U, nF, nU)) return rc;
    s->adaptive_tau = tau;                                                  // :486
    if(info)
    {
        info->Phi_iters = i; info->T_iters = i;                             // :488-492
        info->Phi_converged = converged; info->T_converged = converged;
        info->Phi_error = eps_F; info->T_error = eps_U;
        info->next_tau = tau;
    }
    *dt_out = used_tau;
    return PF_OK;
}

// -------------------------------------------------------------------------------------------------------
// Semi-implicit solver
// -------------------------------------------------------------------------------------------------------
static inline int cg_blocks(const pf_solver* s, int nx, int ny)
{
    int segs = (nx + CG_THREADS - 1) / CG_THREADS;
    long long items = (long long) segs * ny;
    long long want = std::min<long long>(items, (long long) s->sm_count * s->si_blocks_per_sm);
    return (int) std::max<long long>(1, std::min<long long>(want, CG_MAX_BLOCKS));
}

// conjugate_gradient_solve: reference simulation.cu:596-690. All scalars stay on the device; `which` selects the CgState slot.
template <typename Real>
static int cg_solve_device(pf_solver* s, const CgMatrix<Real>& A, Real* x, const Real* b, const Real* x0, Real tolerance, int max_iters,
                           Real* r, Real* pv, Real* Ap, int slot)
{
    const int nx = A.nx, ny = A.ny;
    const size_t N = (size_t) nx * (size_t) ny;
    const int blocks = cg_blocks(s, nx, ny);
    Real* P = (Real*) s->partials;
    Real* rrP[2] = {P, P + CG_MAX_BLOCKS};
    Real* pApP = P + 2 * CG_MAX_BLOCKS;
    CgState* st = s->d_cg + slot;
    const Real epsilon = (Real) 1.0e-12;                                    // simulation.cu:882
    const Real sst = tolerance * tolerance * (Real)(int) N;                 // :608
    k_cg_init<Real><<<blocks, CG_THREADS, 0, s->stream>>>(A, x0, b, x, r, pv, rrP[0], st, max_iters); PF_LAUNCH_CHECK(s);
    for(int it = 0; it < max_iters; it++)
    {
        k_cg_spmv<Real><<<blocks, CG_THREADS, 0, s->stream>>>(A, pv, Ap, pApP, st); PF_LAUNCH_CHECK(s);
        k_cg_update<Real><<<blocks, CG_THREADS, 0, s->stream>>>(N, epsilon, x, r, pv, Ap, rrP[it & 1], pApP, rrP[(it + 1) & 1], blocks, st); PF_LAUNCH_CHECK(s);
        k_cg_pupdate<Real><<<blocks, CG_THREADS, 0, s->stream>>>(N, epsilon, sst, r, pv, rrP[it & 1], rrP[(it + 1) & 1], blocks, st, it); PF_LAUNCH_CHECK(s);
    }
    if(max_iters <= 0) { k_cg_finalize<Real><<<1, CG_THREADS, 0, s->stream>>>(rrP[0], blocks, st); PF_LAUNCH_CHECK(s); }
    return PF_OK;
}

template <typename Real>
static CgMatrix<Real> make_matrix(const pf_params& p, int which, const Real* factor)
{
    CgMatrix<Real> A; memset(&A, 0, sizeof A);
    const Real dt = (Real) p.dt, dx = (Real) p.L0 / p.nx, dy = (Real) p.L0 / p.ny;
    A.which = which; A.factor = factor; A.nx = p.nx; A.ny = p.ny; A.y_begin = 0; A.y_end = p.ny;
    if(which == 0)
    {
        A.C_minus_one = 2 * dt / (dx * dx) + 2 * dt / (dy * dy);          // simulation.cu:774-776
        A.X = -dt / (dx * dx);
        A.Y = -dt / (dy * dy);
        A.boundary = p.Phi_boundary;
    }
    else
    {
        A.C = 1 + 2 * dt / (dx * dx) + 2 * dt / (dy * dy);                // simulation.cu:783-787
        A.R = -dt / (dx * dx); A.Lc = -dt / (dx * dx);
        A.Uc = -dt / (dy * dy); A.D = -dt / (dy * dy);
        A.boundary = p.T_boundary;
    }
    return A;
}

// The whole step as ONE cooperative launch (pf_si_coop.cuh) when its grid can be co-resident; *used = 0 otherwise.
template <typename Real, bool ANISO>
static int semi_implicit_step_coop_inst(pf_solver* s, const pf_params& p, const Real* F, const Real* U, const Real* U_base, Real* nF, Real* nU, int* used)
{
    *used = 0;
    const int slot = (sizeof(Real) == 8 ? 0 : 2) + (ANISO ? 1 : 0);
    auto kern = k_si_step_coop<Real, ANISO>;
    if(s->coop_blocks_per_sm[slot] < 0)
    {
        int coop = 0, per_sm = 0;
        PF_CUDA(s, cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, s->device));
        if(coop) PF_CUDA(s, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, CG_THREADS, 0));
        s->coop_blocks_per_sm[slot] = coop ? per_sm : 0;
    }
    const int blocks = cg_blocks(s, s->nx, s->ny);
    if(s->force_per_kernel_cg || (long long) s->coop_blocks_per_sm[slot] * s->sm_count < blocks) return PF_OK;

    void* q[5];
    for(int i = 0; i < 5; i++) if(int rc = get_scratch(s, 4 + i, &q[i])) return rc;     // the per-kernel path's factor, r, p, Ap slots + the second p buffer
    Real* P = (Real*) s->partials;
    SiCoopArgs<Real> a; memset(&a, 0, sizeof a);
    a.F = F; a.U = U; a.U_base = U_base; a.nF = nF; a.nU = nU;
    a.factor = (Real*) q[0]; a.r = (Real*) q[1]; a.p = (Real*) q[2]; a.Ap = (Real*) q[3]; a.p2 = (Real*) q[4];
    a.rr0 = P; a.rr1 = P + CG_MAX_BLOCKS; a.pAp = P + 2 * CG_MAX_BLOCKS;
    a.st = s->d_cg;
    a.A_F = make_matrix<Real>(p, 0, a.factor);
    a.A_U = make_matrix<Real>(p, 1, nullptr);
    a.k = make_consts<Real>(p);
    a.nx = s->nx; a.ny = s->ny; a.F_bound = p.Phi_boundary; a.U_bound = p.T_boundary;
    const size_t N = s->field_elems();
    a.sst_F = (Real) p.Phi_tolerance * (Real) p.Phi_tolerance * (Real)(int) N;         // simulation.cu:608
    a.sst_U = (Real) p.T_tolerance * (Real) p.T_tolerance * (Real)(int) N;
    a.max_F = p.Phi_max_iters; a.max_U = p.T_max_iters;
    a.L = (Real) p.L; a.dt_omg = (Real) p.dt * (1 - (Real) p.gamma);
    if(!s->plain_grid_sync)
    {
        a.ll = s->ll_slots; a.ll_epoch = s->ll_epoch;
        const unsigned used_epochs = 2u * (1u + 2u * (unsigned) std::max(std::max(a.max_F, a.max_U), 0)) + 2u;   // one per GridSum call of the launch, at most
        s->ll_epoch += used_epochs;
        if(s->ll_epoch < used_epochs + 1u) s->ll_epoch = 1;                 // 32-bit wrap: start over (the slots then hold epochs far away)
    }
    void* args[] = {&a};
    PF_CUDA(s, cudaLaunchCooperativeKernel((const void*) kern, dim3((unsigned) blocks), dim3(CG_THREADS), args, 0, s->stream));
    PF_LAUNCH_CHECK(s);
    *used = 1;
    return PF_OK;
}

template <typename Real>
static int semi_implicit_step_coop(pf_solver* s, const pf_params& p, const Real* F, const Real* U, const Real* U_base, Real* nF, Real* nU, int* used)
{
    if(p.S != 0) return semi_implicit_step_coop_inst<Real, true>(s, p, F, U, U_base, nF, nU, used);
    return semi_implicit_step_coop_inst<Real, false>(s, p, F, U, U_base, nF, nU, used);
}

// semi_implicit_solver_step_based: reference simulation.cu:732-926 (without the do_debug back-substitution).
template <typename Real>
static int semi_implicit_step_based(pf_solver* s, const pf_params& p, const Real* F, const Real* U, const Real* U_base, Real* nF, Real* nU)
{
    if(!s->is_full_grid()) return fail(s, PF_ERR_UNSUPPORTED, "the semi-implicit solver is single-GPU only in this version");
    {
        int used = 0;
        if(int rc = semi_implicit_step_coop<Real>(s, p, F, U, U_base, nF, nU, &used)) return rc;
        if(used) return PF_OK;
    }
    void* q[6];
    for(int i = 0; i < 6; i++) if(int rc = get_scratch(s, 2 + i, &q[i])) return rc;     // slots 0,1 are the corrector's temporaries
    Real *b_F = (Real*) q[0], *b_U = (Real*) q[1], *aniso = (Real*) q[2], *r = (Real*) q[3], *pv = (Real*) q[4], *Ap = (Real*) q[5];
    KConsts<Real> k = make_consts<Real>(p);
    dim3 block(32, 8), grid((unsigned)((s->nx + 31) / 32), (unsigned)((s->ny + 7) / 8));
    if(p.S != 0) k_si_prepare<Real, true><<<grid, block, 0, s->stream>>>(F, U, b_F, aniso, s->nx, s->ny, p.Phi_boundary, p.T_boundary, k);
    else         k_si_prepare<Real, false><<<grid, block, 0, s->stream>>>(F, U, b_F, aniso, s->nx, s->ny, p.Phi_boundary, p.T_boundary, k);
    PF_LAUNCH_CHECK(s);

    CgMatrix<Real> A_F = make_matrix<Real>(p, 0, aniso);
    if(int rc = cg_solve_device<Real>(s, A_F, nF, b_F, F, (Real) p.Phi_tolerance, p.Phi_max_iters, r, pv, Ap, 0)) return rc;   // :881-888

    const size_t N = s->field_elems();
    const Real dt = (Real) p.dt, gamma = (Real) p.gamma;
    k_si_bU<Real><<<elementwise_blocks(s, N, 256), 256, 0, s->stream>>>(N, (Real) p.L, dt * (1 - gamma), U_base, F, nF, b_U);   // :893-899
    PF_LAUNCH_CHECK(s);

    CgMatrix<Real> A_U = make_matrix<Real>(p, 1, nullptr);
    if(int rc = cg_solve_device<Real>(s, A_U, nU, b_U, U, (Real) p.T_tolerance, p.T_max_iters, r, pv, Ap, 1)) return rc;       // :901-906
    return PF_OK;
}

template <typename Real>
static int stats_delta_device(pf_solver* s, const Real* a, const Real* b, pf_delta_stats* out)
{
    const size_t N = s->field_elems();
    int blocks = std::min<int>(CG_MAX_BLOCKS, elementwise_blocks(s, N, CG_THREADS));
    k_stats_delta<Real><<<blocks, CG_THREADS, 0, s->stream>>>(N, a, b, s->d_stats); PF_LAUNCH_CHECK(s);
    PF_CUDA(s, cudaMemcpyAsync(s->h_stats, s->d_stats, blocks * sizeof(StatsPartial), cudaMemcpyDeviceToHost, s->stream));
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    Real sum = 0, L1 = 0, L2 = 0, mn = (Real) INFINITY, mx = -(Real) INFINITY;   // fixed order: bit-stable
    for(int i = 0; i < blocks; i++)
    {
        sum += (Real) s->h_stats[i].sum; L1 += (Real) s->h_stats[i].L1; L2 += (Real) s->h_stats[i].L2;
        mn = std::min(mn, (Real) s->h_stats[i].mn); mx = std::max(mx, (Real) s->h_stats[i].mx);
    }
    out->sum = (double) sum; out->L1 = (double) L1; out->L2 = (double)(Real) sqrt((double) L2);
    out->min = (double) mn; out->max = (double) mx;
    return PF_OK;
}

// semi_implicit_and_euler_solver_step_corrector: reference simulation.cu:928-1008.
template <typename Real>
static int step_corrector(pf_solver* s, const pf_params& p, const Real* F, const Real* U, Real* nF, Real* nU, pf_step_info* info)
{
    size_t max_iters = (size_t) std::max(0, p.corrector_max_iters);
    if(!p.do_corrector_loop) max_iters = 0;
    if(max_iters == 0 && p.do_stats_step_residual) max_iters = 1;

    Real* steps_F[2]; Real* steps_U[2];
    Real *tF = nullptr, *tU = nullptr;
    const bool need_temp = max_iters > 0 || (s->last_num_steps % 2) != 0;
    if(need_temp)
    {
        if(!s->is_full_grid()) return fail(s, PF_ERR_UNSUPPORTED, "the corrector loop is single-GPU only");
        void* q; if(int rc = get_scratch(s, 10, &q)) return rc; tF = (Real*) q;
        if(int rc = get_scratch(s, 11, &q)) return rc; tU = (Real*) q;
    }
    if(s->last_num_steps % 2 == 0) { steps_F[0] = nF; steps_U[0] = nU; steps_F[1] = tF; steps_U[1] = tU; }
    else                           { steps_F[1] = nF; steps_U[1] = nU; steps_F[0] = tF; steps_U[0] = tU; }

    const bool euler = p.solver == PF_SOLVER_EXPLICIT_EULER;
    if(euler) { if(int rc = euler_step_based<Real>(s, p, F, U, U, steps_F[0], steps_U[0], info)) return rc; }
    else      { if(int rc = semi_implicit_step_based<Real>(s, p, F, U, U, steps_F[0], steps_U[0])) return rc; }
    for(size_t k = 0; k < max_iters; k++)
    {
        int curr = (int)(k % 2), next = (int)((k + 1) % 2);
        if(euler) { if(int rc = euler_step_based<Real>(s, p, F, steps_U[curr], U, steps_F[next], steps_U[next], info)) return rc; }
        else      { if(int rc = semi_implicit_step_based<Real>(s, p, F, steps_U[curr], U, steps_F[next], steps_U[next])) return rc; }
        if(p.do_stats_step_residual && info && k < PF_MAX_STEP_RESIDUALS)
        {
            if(int rc = stats_delta_device<Real>(s, steps_F[curr], steps_F[next], &info->step_res[k])) return rc;
            info->step_res_count = (int) max_iters;
        }
        s->last_num_steps = (int) k;                                        // :995
    }
    int final_step = ((s->last_num_steps % 2) + 2) % 2;                     // :999
    if(steps_F[final_step] != nF)
    {
        PF_CUDA(s, cudaMemcpyAsync(nF, steps_F[final_step], s->field_bytes(), cudaMemcpyDeviceToDevice, s->stream));
        PF_CUDA(s, cudaMemcpyAsync(nU, steps_U[final_step], s->field_bytes(), cudaMemcpyDeviceToDevice, s->stream));
    }
    return PF_OK;
}

// sim_step: reference simulation.cu:1091-1156. `sync` = return only when the device is done (the reference always does).
template <typename Real>
static int step_impl(pf_solver* s, const pf_params& p, const void* F_, const void* U_, void* nF_, void* nU_, pf_step_info* info, bool sync)
{
    const Real* F = (const Real*) F_; const Real* U = (const Real*) U_; Real* nF = (Real*) nF_; Real* nU = (Real*) nU_;
    pf_step_info local; if(!info) info = &local;
    memset(info, 0, sizeof *info);
    double dt = 0;
    switch(p.solver)
    {
        case PF_SOLVER_SEMI_IMPLICIT:
        case PF_SOLVER_EXPLICIT_EULER:
            if(int rc = step_corrector<Real>(s, p, F, U, nF, nU, info)) return rc;
            dt = p.dt; break;
        case PF_SOLVER_EXPLICIT_RK4:
            if(int rc = rk4_step<Real>(s, p, F, U, nF, nU, info)) return rc;
            dt = p.dt; break;
        case PF_SOLVER_EXPLICIT_RK4_ADAPTIVE:
            if(int rc = rk4_adaptive_step<Real>(s, p, F, U, nF, nU, info, &dt)) return rc;
            break;
        case PF_SOLVER_EXACT:                                               // exact_solver_step, :1069-1089, :1118-1121
        {
            if(!s->is_full_grid()) return fail(s, PF_ERR_UNSUPPORTED, "solver = exact is single-GPU only");
            dim3 grid((unsigned)((s->nx + 255) / 256), (unsigned) s->ny);
            k_exact_fields<Real><<<grid, 256, 0, s->stream>>>(s->nx, s->ny, (Real) p.L0 / p.nx, (Real) p.L0 / p.ny, (Real) p.L0, (Real) p.time, nF, nU);
            PF_LAUNCH_CHECK(s);
            dt = p.dt;
            break;
        }
        default:
            return fail(s, PF_ERR_INVALID, "unknown solver %d", p.solver);
    }
    if(p.do_stats)
    {
        if(int rc = stats_delta_device<Real>(s, F, nF, &info->Phi_delta)) return rc;       // :1128-1129
        if(int rc = stats_delta_device<Real>(s, U, nU, &info->T_delta)) return rc;
    }
    if(p.solver == PF_SOLVER_SEMI_IMPLICIT && sync)
    {
        PF_CUDA(s, cudaMemcpyAsync(s->h_cg, s->d_cg, 2 * sizeof(CgState), cudaMemcpyDeviceToHost, s->stream));
        PF_CUDA(s, cudaStreamSynchronize(s->stream));
        const double N = (double) s->nx * (double) s->ny;
        info->Phi_iters = s->h_cg[0].iters; info->T_iters = s->h_cg[1].iters;
        info->Phi_converged = s->h_cg[0].iters != p.Phi_max_iters; info->T_converged = s->h_cg[1].iters != p.T_max_iters;
        info->Phi_error = (double)(Real) sqrt((double)(Real)((Real) s->h_cg[0].rr / (Real) N));
        info->T_error = (double)(Real) sqrt((double)(Real)((Real) s->h_cg[1].rr / (Real) N));
    }
    else if(sync)
        PF_CUDA(s, cudaStreamSynchronize(s->stream));                       // :1153
    info->dt_advanced = dt;
    return PF_OK;
}

static int check_step_args(pf_solver* s, const pf_params* p, const void* F, const void* U, const void* nF, const void* nU)
{
    if(!p || !F || !U || !nF || !nU) return fail(s, PF_ERR_INVALID, "pf_step: NULL argument");
    if(p->nx != s->nx || p->ny != s->ny)                                    // simulation.cu:1094-1095
        return fail(s, PF_ERR_INVALID, "Bad sizes provided: params %dx%d, solver %dx%d", p->nx, p->ny, s->nx, s->ny);
    if(p->do_exact && !s->is_full_grid()) return fail(s, PF_ERR_UNSUPPORTED, "do_exact is single-GPU only");
    if(F == nF || U == nU) return fail(s, PF_ERR_INVALID, "pf_step: in-place stepping is not supported");
    for(int b : {p->Phi_boundary, p->T_boundary})
        if(b < PF_BOUNDARY_PERIODIC || b > PF_BOUNDARY_NEUMANN_ZERO) return fail(s, PF_ERR_INVALID, "bad boundary type %d", b);
    if(!s->is_full_grid() && s->slab.ghost < pf_required_ghost(p->solver))
        return fail(s, PF_ERR_INVALID, "slab ghost depth %d < %d required by solver %d", s->slab.ghost, pf_required_ghost(p->solver), p->solver);
    return PF_OK;
}

extern "C" {

int pf_step(pf_solver* s, const pf_params* params, const void* F, const void* U, void* next_F, void* next_U, pf_step_info* info)
{
    if(int rc = bind(s)) return rc;
    if(int rc = check_step_args(s, params, F, U, next_F, next_U)) return rc;
    return s->dtype == PF_F64 ? step_impl<double>(s, *params, F, U, next_F, next_U, info, true)
                              : step_impl<float>(s, *params, F, U, next_F, next_U, info, true);
}

int pf_step_async(pf_solver* s, const pf_params* params, const void* F, const void* U, void* next_F, void* next_U)
{
    if(int rc = bind(s)) return rc;
    if(int rc = check_step_args(s, params, F, U, next_F, next_U)) return rc;
    if(params->solver == PF_SOLVER_EXPLICIT_RK4_ADAPTIVE) return fail(s, PF_ERR_UNSUPPORTED, "pf_step_async: the adaptive solver answers dt on the host");
    if(params->do_stats || params->do_stats_step_residual) return fail(s, PF_ERR_UNSUPPORTED, "pf_step_async: statistics need the host");
    return s->dtype == PF_F64 ? step_impl<double>(s, *params, F, U, next_F, next_U, nullptr, false)
                              : step_impl<float>(s, *params, F, U, next_F, next_U, nullptr, false);
}

// pf_run's batches of single-launch steps (explicit Euler on the tile kernel; explicit-rk4 where the multi-step launch does not
// apply): RUN_GRAPH_STEPS steps, the buffer ping-pong baked in (an even count: the replay ends where it began), captured once
// per {buffers, parameters} and replayed. At 128^2 - 512^2 a step is a 2 - 3 us kernel behind a 2 - 4 us launch; the replay takes
// the launches off the host. Nothing of a fused-eligible step depends on iter / time, so one capture serves the whole run.
enum { RUN_GRAPH_STEPS = 64 };
static int run_graph_prepare(pf_solver* s, const pf_params& p, void* cF, void* cU, void* nF, void* nU)
{
    pf_solver::RunGraphKey key; memset(&key, 0, sizeof key);
    key.F = cF; key.U = cU; key.nF = nF; key.nU = nU; key.p = p; key.p.iter = 0; key.p.time = 0;
    if(s->run_graph && memcmp(&key, &s->run_graph_key, sizeof key) == 0) return PF_OK;
    if(s->run_graph) { cudaGraphExecDestroy(s->run_graph); s->run_graph = nullptr; }
    const int64_t launches0 = s->launches;
    cudaGraph_t g = nullptr;
    PF_CUDA(s, cudaStreamBeginCapture(s->stream, cudaStreamCaptureModeThreadLocal));
    int rc = PF_OK;
    void *a = cF, *b = cU, *c = nF, *d = nU;
    for(int k = 0; k < RUN_GRAPH_STEPS && rc == PF_OK; k++)
    {
        pf_step_info info;
        rc = s->dtype == PF_F64 ? step_impl<double>(s, p, a, b, c, d, &info, false) : step_impl<float>(s, p, a, b, c, d, &info, false);
        std::swap(a, c); std::swap(b, d);
    }
    cudaError_t e = cudaStreamEndCapture(s->stream, &g);
    const int64_t per_replay = s->launches - launches0;
    s->launches = launches0;                                                // captured, not launched
    if(rc == PF_OK && e == cudaSuccess && g && per_replay == RUN_GRAPH_STEPS) e = cudaGraphInstantiate(&s->run_graph, g, 0);
    else if(e == cudaSuccess) e = cudaErrorUnknown;
    if(g) cudaGraphDestroy(g);
    if(e != cudaSuccess || !s->run_graph)
    {
        cudaGetLastError();                                                 // this handle goes on without graphs
        s->run_graph = nullptr; s->run_graphs_ok = false;
        return PF_OK;
    }
    memcpy(&s->run_graph_key, &key, sizeof key);
    return PF_OK;
}

int pf_run(pf_solver* s, const pf_params* params, void* F, void* U, void* F_tmp, void* U_tmp, int64_t n_steps, double stop_time,
           void** F_out, void** U_out, int64_t* steps_done, double* t_advanced, pf_step_info* last_info)
{
    return pf_run_until(s, params, F, U, F_tmp, U_tmp, n_steps, stop_time, 0.0, F_out, U_out, steps_done, t_advanced, nullptr, nullptr, last_info);
}

int pf_run_until(pf_solver* s, const pf_params* params, void* F, void* U, void* F_tmp, void* U_tmp, int64_t n_steps, double stop_time,
                 double pause_time, void** F_out, void** U_out, int64_t* steps_done, double* t_advanced, double* t_end, double* t_last_begin,
                 pf_step_info* last_info)
{
    if(int rc = bind(s)) return rc;
    if(int rc = check_step_args(s, params, F, U, F_tmp, U_tmp)) return rc;
    pf_params p = *params;
    void *cF = F, *cU = U, *nF = F_tmp, *nU = U_tmp;
    double t = params->time, t_before_last = params->time;
    int64_t i = 0;
    // The delta statistics (simulation.cu:1126-1142) of every step but the last would be thrown away (one pf_step_info comes
    // back): the steps run without them, on the fast paths, and the last step's statistics are formed after the loop from the
    // two buffers of the ping-pong - the state before the last step is still in the other one.
    const bool want_stats = p.do_stats != 0 && !p.do_stats_step_residual;
    if(want_stats) p.do_stats = 0;
    // fixed-dt solvers without statistics never need the host between steps: enqueue everything, synchronise once
    const bool async_ok = p.solver != PF_SOLVER_EXPLICIT_RK4_ADAPTIVE && !p.do_stats && !p.do_stats_step_residual;
    // the fixed-step explicit solvers on the single-kernel path: ONE launch per batch of steps (k_stream_step, StreamRun)
    const bool multi_ok = async_ok && (p.solver == PF_SOLVER_EXPLICIT_EULER || p.solver == PF_SOLVER_EXPLICIT_RK4) && !p.do_corrector_loop &&
                          !p.do_corrector_guess && fused_eligible(s, p) && (s->last_num_steps % 2) == 0 && s->is_full_grid();
    while(i < n_steps)
    {
        if(stop_time > 0 && stop_time - t < 1e-16) break;                  // main.cpp:525
        if(pause_time > 0 && i > 0 && t >= pause_time) break;              // the host's snapshot tests (main.cpp:507-520) would fire: hand back
        p.iter = params->iter + i; p.time = t;
        int64_t batch = 1;
        if(multi_ok)
        {
            // how many steps until the stop time (the same additions of dt the loop below would make), at most 2^20 per launch
            double tt = t;
            batch = 0;
            while(i + batch < n_steps && batch < (1 << 20) && !(stop_time > 0 && stop_time - tt < 1e-16) &&
                  !(pause_time > 0 && i + batch > 0 && tt >= pause_time)) { tt += p.dt; batch++; }
        }
        pf_step_info info;
        int64_t graph_steps = 0;
        if(batch > 1)
        {
            s->run_steps = (int) batch;
            int rc = s->dtype == PF_F64 ? step_impl<double>(s, p, cF, cU, nF, nU, &info, false) : step_impl<float>(s, p, cF, cU, nF, nU, &info, false);
            s->run_steps = 1;
            if(rc) return rc;
            if(s->run_used)
            {
                for(int64_t k = 0; k < batch; k++) { t_before_last = t; t += info.dt_advanced; }   // the same floating-point additions as step by step
                i += batch;
                if(batch & 1) { std::swap(cF, nF); std::swap(cU, nU); }
                if(last_info) *last_info = info;
                continue;
            }
            // the grid does not fit one wave (or the device cannot launch cooperatively): that was a plain single step
            if(s->run_graphs_ok) graph_steps = (batch - 1) / RUN_GRAPH_STEPS * RUN_GRAPH_STEPS;
        }
        else
        {
            const bool last = (i + 1 == n_steps) || (stop_time > 0 && stop_time - (t + p.dt) < 1e-16 && async_ok);
            int rc = s->dtype == PF_F64 ? step_impl<double>(s, p, cF, cU, nF, nU, &info, !async_ok || last)
                                        : step_impl<float>(s, p, cF, cU, nF, nU, &info, !async_ok || last);
            if(rc) return rc;
        }
        t_before_last = t;
        t += info.dt_advanced;
        if(last_info) *last_info = info;
        std::swap(cF, nF); std::swap(cU, nU);
        i++;
        if(graph_steps > 0)
        {
            if(int rc = run_graph_prepare(s, p, cF, cU, nF, nU)) return rc;
            if(s->run_graph)
            {
                for(int64_t k = 0; k < graph_steps; k += RUN_GRAPH_STEPS)
                {
                    PF_CUDA(s, cudaGraphLaunch(s->run_graph, s->stream));
                    s->launches += RUN_GRAPH_STEPS;
                }
                for(int64_t k = 0; k < graph_steps; k++) { t_before_last = t; t += info.dt_advanced; }   // the same floating-point additions as step by step
                i += graph_steps;                                                  // an even number of steps: the buffers are where they were
            }
        }
    }
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    if(s->run_progress)
    {
        unsigned stalled = 0;
        PF_CUDA(s, cudaMemcpy(&stalled, s->run_progress + s->run_progress_units, sizeof stalled, cudaMemcpyDeviceToHost));
        if(stalled) { cudaMemset(s->run_progress + s->run_progress_units, 0, sizeof stalled); return fail(s, PF_ERR_CUDA, "pf_run: a multi-step launch stalled (its CTAs were not co-resident)"); }
    }
    if(want_stats && i > 0 && last_info)
    {
        int rc;
        if(s->dtype == PF_F64) { rc = stats_delta_device<double>(s, (const double*) nF, (const double*) cF, &last_info->Phi_delta);
                                 if(!rc) rc = stats_delta_device<double>(s, (const double*) nU, (const double*) cU, &last_info->T_delta); }
        else                   { rc = stats_delta_device<float>(s, (const float*) nF, (const float*) cF, &last_info->Phi_delta);
                                 if(!rc) rc = stats_delta_device<float>(s, (const float*) nU, (const float*) cU, &last_info->T_delta); }
        if(rc) return rc;
    }
    if(F_out) *F_out = cF;
    if(U_out) *U_out = cU;
    if(steps_done) *steps_done = i;
    if(t_advanced) *t_advanced = t - params->time;
    if(t_last_begin) *t_last_begin = t_before_last;                         // the time the last step started from (its stats row, main.cpp:192-234)
    if(t_end) *t_end = t;                                                   // params->time + dt + dt + ...: the host loop's own additions
    return PF_OK;
}

// Scoped row window + stream selection for the *_rows_async entry points.
struct RowWindow {
    pf_solver* s;
    RowWindow(pf_solver* s_, int y0, int y1, int on_aux) : s(s_) { s->win_y0 = y0; s->win_y1 = y1; s->launch_stream = on_aux ? s->aux_stream : s->stream; }
    ~RowWindow() { s->win_y0 = s->win_y1 = 0; s->launch_stream = nullptr; }
};

int pf_step_rows_async(pf_solver* s, const pf_params* params, const void* F, const void* U, void* next_F, void* next_U,
                       int y_begin, int y_end, int on_aux_stream)
{
    if(int rc = bind(s)) return rc;
    if(int rc = check_step_args(s, params, F, U, next_F, next_U)) return rc;
    if(params->solver != PF_SOLVER_EXPLICIT_EULER && params->solver != PF_SOLVER_EXPLICIT_RK4)
        return fail(s, PF_ERR_UNSUPPORTED, "pf_step_rows_async: explicit / explicit-rk4 only (use pf_rkm_attempt_rows_async for the adaptive solver)");
    if(params->do_corrector_loop || params->do_stats || params->do_stats_step_residual || !fused_eligible(s, *params) || (s->last_num_steps % 2) != 0)
        return fail(s, PF_ERR_UNSUPPORTED, "pf_step_rows_async: needs the fused single-kernel path (equal boundary types, no corrector loop, no statistics)");
    if(y_end <= y_begin) return PF_OK;
    RowWindow w(s, y_begin, y_end, on_aux_stream);
    return s->dtype == PF_F64 ? step_impl<double>(s, *params, F, U, next_F, next_U, nullptr, false)
                              : step_impl<float>(s, *params, F, U, next_F, next_U, nullptr, false);
}

// sim_modify(UPLOAD) x2 + sim_step + sim_modify(DOWNLOAD) x2 as ONE call (main.cpp:164-168, :553, :857): the host hands
// in y(n) and gets y(n+1) back. For the fixed-step explicit solvers on the single-kernel path the three legs are
// pipelined over row bands: band b+1 uploads while band b computes while band b-1 downloads (a band needs `ghost` rows
// of its neighbours, so its kernel waits for the upload of band b+1). PCIe is full duplex, so the step costs about one
// direction's copy time instead of upload + compute + download. Everything else takes the plain sequence.
int pf_step_host(pf_solver* s, const pf_params* params, const void* host_F, const void* host_U, void* host_next_F, void* host_next_U,
                 int host_dtype, void* F, void* U, void* next_F, void* next_U, pf_step_info* info)
{
    if(int rc = bind(s)) return rc;
    if(!host_F || !host_U || !host_next_F || !host_next_U) return fail(s, PF_ERR_INVALID, "pf_step_host: NULL host pointer");
    if(int rc = check_step_args(s, params, F, U, next_F, next_U)) return rc;
    const size_t count = (size_t) s->nx * (size_t) s->slab.ny_local;
    const size_t first = (size_t) s->ghost_lo * (size_t) s->nx;            // element offset of the first owned row
    const size_t rs = s->real_size();
    char *dF = (char*) F + first * rs, *dU = (char*) U + first * rs, *dnF = (char*) next_F + first * rs, *dnU = (char*) next_U + first * rs;
    const bool pipelined = host_dtype == s->dtype && s->is_full_grid() &&
        (params->solver == PF_SOLVER_EXPLICIT_EULER || params->solver == PF_SOLVER_EXPLICIT_RK4) &&
        !params->do_corrector_loop && !params->do_stats_step_residual && fused_eligible(s, *params) &&
        (s->last_num_steps % 2) == 0 && s->ny >= 64;
    if(!pipelined)
    {
        if(int rc = pf_upload(s, dF, host_F, count, host_dtype)) return rc;
        if(int rc = pf_upload(s, dU, host_U, count, host_dtype)) return rc;
        if(int rc = pf_step(s, params, F, U, next_F, next_U, info)) return rc;
        if(int rc = pf_download(s, dnF, host_next_F, count, host_dtype)) return rc;
        return pf_download(s, dnU, host_next_U, count, host_dtype);
    }
    if(!s->h2d_stream)
    {
        PF_CUDA(s, cudaStreamCreateWithFlags(&s->h2d_stream, cudaStreamNonBlocking));
        PF_CUDA(s, cudaStreamCreateWithFlags(&s->d2h_stream, cudaStreamNonBlocking));
        PF_CUDA(s, cudaEventCreateWithFlags(&s->ev_host_begin, cudaEventDisableTiming));
        for(int i = 0; i < pf_solver::HOST_BANDS_MAX; i++)
        {
            PF_CUDA(s, cudaEventCreateWithFlags(&s->ev_up[i], cudaEventDisableTiming));
            PF_CUDA(s, cudaEventCreateWithFlags(&s->ev_done[i], cudaEventDisableTiming));
        }
    }
    // bands of >= 64 rows, at most HOST_BANDS_MAX: more bands shorten the pipeline fill and drain but every copy and launch costs
    // enqueue time (numbers below); the two copy directions alone need 1.4 ms when overlapped (2 x 32 MiB each way at 2048^2 fp64).
    const int ny = s->ny;
    const int nb = std::max(1, std::min<int>({s->host_bands, (int) pf_solver::HOST_BANDS_MAX, ny / 64}));
    const int rows_per = (ny + nb - 1) / nb;
    const size_t row_bytes = (size_t) s->nx * rs;
    // the copies must not overtake work already queued on the solver's stream that still reads or writes these fields
    PF_CUDA(s, cudaEventRecord(s->ev_host_begin, s->stream));
    PF_CUDA(s, cudaStreamWaitEvent(s->h2d_stream, s->ev_host_begin, 0));
    // Periodic in y: band 0 also reads the LAST `ghost` rows of the grid (wrapped), which belong to the last band. They go up
    // first, ahead of band 0, so that every band's inputs are on the device once the upload of the band after it is.
    const int wrap_rows = params->Phi_boundary == PF_BOUNDARY_PERIODIC ? std::min(ny, pf_required_ghost(params->solver)) : 0;
    if(wrap_rows)
    {
        const size_t off = (size_t)(ny - wrap_rows) * row_bytes, bytes = (size_t) wrap_rows * row_bytes;
        PF_CUDA(s, cudaMemcpyAsync(dF + off, (const char*) host_F + off, bytes, cudaMemcpyHostToDevice, s->h2d_stream));
        PF_CUDA(s, cudaMemcpyAsync(dU + off, (const char*) host_U + off, bytes, cudaMemcpyHostToDevice, s->h2d_stream));
    }
    // Upload piece b ends `ghost` rows beyond band b, so that band b's inputs (its rows and `ghost` rows on either side) are on the
    // device as soon as ITS piece is: compute b never waits for the whole next piece, and the downloads chain right behind the uploads.
    // (Measured on the same box, 2048^2 fp64 explicit-rk4: 1.96 -> 1.89 ms per call at 8 bands; 4 bands 1.84, 16 bands 2.14: every band
    // costs ~35 us of enqueue and copy set-up. Phi and T on two copy streams per direction: slower, 2.3 ms.)
    const int ghost = pf_required_ghost(params->solver);
    const int ny_up = ny - wrap_rows;                                            // the wrapped tail is already queued
    pf_params band_params = *params; band_params.do_stats = 0;            // the statistics are formed once, after the last band (below)
    int bands = 0, up0 = 0;
    for(int y0 = 0; y0 < ny; y0 += rows_per, bands++)
    {
        const int y1 = std::min(ny, y0 + rows_per);
        const int up1 = std::max(up0, std::min(ny_up, y1 + ghost));
        const size_t off = (size_t) up0 * row_bytes, bytes = (size_t)(up1 - up0) * row_bytes;
        if(bytes)
        {
            PF_CUDA(s, cudaMemcpyAsync(dF + off, (const char*) host_F + off, bytes, cudaMemcpyHostToDevice, s->h2d_stream));
            PF_CUDA(s, cudaMemcpyAsync(dU + off, (const char*) host_U + off, bytes, cudaMemcpyHostToDevice, s->h2d_stream));
        }
        PF_CUDA(s, cudaEventRecord(s->ev_up[bands], s->h2d_stream));
        up0 = up1;
    }
    for(int b = 0; b < bands; b++)
    {
        const int y0 = b * rows_per, y1 = std::min(ny, y0 + rows_per);
        PF_CUDA(s, cudaStreamWaitEvent(s->stream, s->ev_up[b], 0));
        {
            RowWindow w(s, y0, y1, 0);
            int rc = s->dtype == PF_F64 ? step_impl<double>(s, band_params, F, U, next_F, next_U, nullptr, false)
                                        : step_impl<float>(s, band_params, F, U, next_F, next_U, nullptr, false);
            if(rc) return rc;
        }
        PF_CUDA(s, cudaEventRecord(s->ev_done[b], s->stream));
        PF_CUDA(s, cudaStreamWaitEvent(s->d2h_stream, s->ev_done[b], 0));
        const size_t off = (size_t) y0 * row_bytes, bytes = (size_t)(y1 - y0) * row_bytes;
        PF_CUDA(s, cudaMemcpyAsync((char*) host_next_F + off, dnF + off, bytes, cudaMemcpyDeviceToHost, s->d2h_stream));
        PF_CUDA(s, cudaMemcpyAsync((char*) host_next_U + off, dnU + off, bytes, cudaMemcpyDeviceToHost, s->d2h_stream));
    }
    PF_CUDA(s, cudaStreamSynchronize(s->d2h_stream));
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    if(info)
    {
        memset(info, 0, sizeof *info);
        info->dt_advanced = params->dt;
        info->Phi_iters = 1; info->T_iters = 1;
        if(params->do_stats)                                                // simulation.cu:1126-1142, on the fields that are now both on the device
        {
            int rc;
            if(s->dtype == PF_F64) { rc = stats_delta_device<double>(s, (const double*) F, (const double*) next_F, &info->Phi_delta);
                                     if(!rc) rc = stats_delta_device<double>(s, (const double*) U, (const double*) next_U, &info->T_delta); }
            else                   { rc = stats_delta_device<float>(s, (const float*) F, (const float*) next_F, &info->Phi_delta);
                                     if(!rc) rc = stats_delta_device<float>(s, (const float*) U, (const float*) next_U, &info->T_delta); }
            if(rc) return rc;
        }
    }
    return PF_OK;
}

int pf_rkm_attempt_rows_async(pf_solver* s, const pf_params* params, double tau, const void* F, const void* U, void* next_F, void* next_U,
                              int y_begin, int y_end, int on_aux_stream)
{
    if(int rc = bind(s)) return rc;
    if(int rc = check_step_args(s, params, F, U, next_F, next_U)) return rc;
    if(!s->rkm_fused) return fail(s, PF_ERR_UNSUPPORTED, "pf_rkm_attempt_rows_async: needs the fused path (equal boundary types)");
    if(y_end <= y_begin) return PF_OK;
    RowWindow w(s, y_begin, y_end, on_aux_stream);
    s->rkm_keep_errors = true;          // the caller zeroes the error words once per attempt with pf_rkm_clear_errors_async
    int rc = s->dtype == PF_F64 ? rkm_attempt<double>(s, *params, tau, (const double*) F, (const double*) U, (double*) next_F, (double*) next_U)
                                : rkm_attempt<float>(s, *params, tau, (const float*) F, (const float*) U, (float*) next_F, (float*) next_U);
    s->rkm_keep_errors = false;
    return rc;
}

int pf_rkm_clear_errors_async(pf_solver* s, int on_aux_stream)
{
    if(int rc = bind(s)) return rc;
    PF_CUDA(s, cudaMemsetAsync(s->d_err, 0, 2 * sizeof(unsigned long long), on_aux_stream ? s->aux_stream : s->stream));
    return PF_OK;
}

int pf_rkm_begin(pf_solver* s, const pf_params* params, const void* F, const void* U, double* tau)
{
    if(int rc = bind(s)) return rc;
    if(!params || !F || !U || !tau) return fail(s, PF_ERR_INVALID, "pf_rkm_begin: NULL argument");
    if(params->nx != s->nx || params->ny != s->ny) return fail(s, PF_ERR_INVALID, "pf_rkm_begin: size mismatch");
    return s->dtype == PF_F64 ? rkm_begin<double>(s, *params, (const double*) F, (const double*) U, tau)
                              : rkm_begin<float>(s, *params, (const float*) F, (const float*) U, tau);
}

int pf_rkm_attempt_async(pf_solver* s, const pf_params* params, double tau, const void* F, const void* U, void* next_F, void* next_U)
{
    if(int rc = bind(s)) return rc;
    if(int rc = check_step_args(s, params, F, U, next_F, next_U)) return rc;
    return s->dtype == PF_F64 ? rkm_attempt<double>(s, *params, tau, (const double*) F, (const double*) U, (double*) next_F, (double*) next_U)
                              : rkm_attempt<float>(s, *params, tau, (const float*) F, (const float*) U, (float*) next_F, (float*) next_U);
}

void* pf_rkm_error_bits(pf_solver* s) { return s ? (void*) s->d_err : nullptr; }

int pf_rkm_decide(pf_solver* s, const pf_params* params, double tau, uint64_t bits_Phi, uint64_t bits_T,
                  int* converged, int* stop, double* next_tau, double* eps_Phi, double* eps_T)
{
    if(!s || !params || !converged || !stop || !next_tau) return fail(s, PF_ERR_INVALID, "pf_rkm_decide: NULL argument");
    double eF = 0, eU = 0;
    if(s->dtype == PF_F64) rkm_decide<double>(*params, tau, bits_Phi, bits_T, converged, stop, next_tau, &eF, &eU);
    else                   rkm_decide<float>(*params, tau, bits_Phi, bits_T, converged, stop, next_tau, &eF, &eU);
    if(eps_Phi) *eps_Phi = eF;
    if(eps_T) *eps_T = eU;
    return PF_OK;
}

int pf_rkm_finish_async(pf_solver* s, const pf_params* params, double used_tau, double next_tau, const void* F, const void* U, void* next_F, void* next_U)
{
    if(int rc = bind(s)) return rc;
    if(int rc = check_step_args(s, params, F, U, next_F, next_U)) return rc;
    int rc = s->dtype == PF_F64 ? rkm_finish<double>(s, used_tau, (const double*) F, (const double*) U, (double*) next_F, (double*) next_U)
                                : rkm_finish<float>(s, used_tau, (const float*) F, (const float*) U, (float*) next_F, (float*) next_U);
    if(rc) return rc;
    s->adaptive_tau = next_tau;                                             // simulation.cu:486
    return PF_OK;
}

int pf_rhs(pf_solver* s, const pf_params* params, int n_states, const double* weights, const void* const* Fs, const void* const* Us, void* out_F, void* out_U)
{
    if(int rc = bind(s)) return rc;
    if(!params || !weights || !Fs || !Us || !out_F || !out_U || n_states < 1 || n_states > 4) return fail(s, PF_ERR_INVALID, "pf_rhs: bad arguments");
    if(!s->is_full_grid()) return fail(s, PF_ERR_UNSUPPORTED, "pf_rhs is single-GPU only");
    if(params->nx != s->nx || params->ny != s->ny) return fail(s, PF_ERR_INVALID, "pf_rhs: size mismatch");
    int rc;
    if(s->dtype == PF_F64)
    {
        double w[4]; for(int i = 0; i < n_states; i++) w[i] = weights[i];
        rc = launch_rhs_general<double>(s, *params, false, n_states, w, (const double* const*) Fs, (const double* const*) Us, (double*) out_F, (double*) out_U);
    }
    else
    {
        float w[4]; for(int i = 0; i < n_states; i++) w[i] = (float) weights[i];
        rc = launch_rhs_general<float>(s, *params, false, n_states, w, (const float* const*) Fs, (const float* const*) Us, (float*) out_F, (float*) out_U);
    }
    if(rc) return rc;
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    return PF_OK;
}

int pf_cg_solve(pf_solver* s, const pf_params* params, int which, const void* aniso, const void* b, const void* x0, void* x,
                double tolerance, int max_iters, int* iters, int* converged, double* error)
{
    if(int rc = bind(s)) return rc;
    if(!params || !b || !x || (which != 0 && which != 1) || (which == 0 && !aniso)) return fail(s, PF_ERR_INVALID, "pf_cg_solve: bad arguments");
    if(!s->is_full_grid()) return fail(s, PF_ERR_UNSUPPORTED, "pf_cg_solve is single-GPU only");
    if(params->nx != s->nx || params->ny != s->ny) return fail(s, PF_ERR_INVALID, "pf_cg_solve: size mismatch");
    void* q[3];
    for(int i = 0; i < 3; i++) if(int rc = get_scratch(s, 5 + i, &q[i])) return rc;
    int rc;
    if(s->dtype == PF_F64)
    {
        CgMatrix<double> A = make_matrix<double>(*params, which, (const double*) aniso);
        rc = cg_solve_device<double>(s, A, (double*) x, (const double*) b, (const double*) x0, tolerance, max_iters, (double*) q[0], (double*) q[1], (double*) q[2], which);
    }
    else
    {
        CgMatrix<float> A = make_matrix<float>(*params, which, (const float*) aniso);
        rc = cg_solve_device<float>(s, A, (float*) x, (const float*) b, (const float*) x0, (float) tolerance, max_iters, (float*) q[0], (float*) q[1], (float*) q[2], which);
    }
    if(rc) return rc;
    PF_CUDA(s, cudaMemcpyAsync(s->h_cg, s->d_cg, 2 * sizeof(CgState), cudaMemcpyDeviceToHost, s->stream));
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    const double N = (double) s->nx * (double) s->ny;
    if(iters) *iters = s->h_cg[which].iters;
    if(converged) *converged = s->h_cg[which].iters != max_iters;
    if(error) *error = s->dtype == PF_F64 ? sqrt(s->h_cg[which].rr / N) : (double) sqrtf((float) s->h_cg[which].rr / (float) N);
    return PF_OK;
}

int pf_debug_values(pf_solver* s, const pf_params* params, const void* F, const void* U, void* grad_F, void* grad_U, void* aniso)
{
    if(int rc = bind(s)) return rc;
    if(!params || !F || !U || !grad_F || !grad_U || !aniso) return fail(s, PF_ERR_INVALID, "pf_debug_values: NULL argument");
    if(!s->is_full_grid()) return fail(s, PF_ERR_UNSUPPORTED, "pf_debug_values is single-GPU only");
    dim3 block(32, 8), grid((unsigned)((s->nx + 31) / 32), (unsigned)((s->ny + 7) / 8));
    if(s->dtype == PF_F64)
        k_debug_values<double><<<grid, block, 0, s->stream>>>((const double*) F, (const double*) U, (double*) grad_F, (double*) grad_U, (double*) aniso,
                                                                s->nx, s->ny, params->Phi_boundary, params->T_boundary, make_consts<double>(*params));
    else
        k_debug_values<float><<<grid, block, 0, s->stream>>>((const float*) F, (const float*) U, (float*) grad_F, (float*) grad_U, (float*) aniso,
                                                               s->nx, s->ny, params->Phi_boundary, params->T_boundary, make_consts<float>(*params));
    PF_LAUNCH_CHECK(s);
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    return PF_OK;
}

int pf_stats_delta(pf_solver* s, const void* a, const void* b, pf_delta_stats* out)
{
    if(int rc = bind(s)) return rc;
    if(!a || !b || !out) return fail(s, PF_ERR_INVALID, "pf_stats_delta: NULL argument");
    return s->dtype == PF_F64 ? stats_delta_device<double>(s, (const double*) a, (const double*) b, out)
                              : stats_delta_device<float>(s, (const float*) a, (const float*) b, out);
}

} // extern "C"

#include "pf_app_host.inl"
#include "pf_si_slab.inl"
#include "pf_halo.inl"

// ---- pf_stats_delta_static: the delta statistics without any allocation ----------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it). The kernel is k_stats_delta, the one pf_stats_delta
// launches; its per-block partials go into the file-scope device array below instead of the solver handle's cudaMalloc'ed buffer.
// Every GPU that loads the library holds the array (PF_STATIC_STATS_MAX_BLOCKS x 40 B = 40 KB) until the process exits.
// In HOST code `g_static_stats` is only a stand-in for the symbol, not a device address: the address is taken once per device,
// with that device current, by cudaGetSymbolAddress, and that pointer is what the kernel gets.
// All calls on a device share the array, which keeps the last call's contents. No word of it has to start at zero: a call
// writes every slot it later reads (blocks 0 .. blocks-1), and the first thing a call does to a slot is overwrite it; the
// slots beyond are never read. Calls that overlap on the GPU (two streams, two host threads) race on the slots.
static __device__ __align__(256) pf::StatsPartial g_static_stats[PF_STATIC_STATS_MAX_BLOCKS];

namespace {
struct StaticStatsDevice { pf::StatsPartial* ptr = nullptr; int sm_count = 0; };
enum { STATIC_STATS_MAX_DEVICES = 64 };
StaticStatsDevice g_static_dev[STATIC_STATS_MAX_DEVICES];
std::mutex g_static_mutex;

// once per device, with that device current
int static_stats_setup(int device, StaticStatsDevice* out)
{
    std::lock_guard<std::mutex> lock(g_static_mutex);
    StaticStatsDevice& d = g_static_dev[device];
    if(!d.ptr)
    {
        pf::StatsPartial* ptr = nullptr; int sms = 0;
        PF_CUDA(nullptr, cudaGetSymbolAddress((void**) &ptr, g_static_stats));
        PF_CUDA(nullptr, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        if(!ptr || sms <= 0) return fail(nullptr, PF_ERR_CUDA, "pf_stats_delta_static: no address for the scratch array on device %d", device);
        d.sm_count = sms; d.ptr = ptr;
    }
    *out = d;
    return PF_OK;
}

template <typename Real>
int stats_delta_static(const StaticStatsDevice& d, const Real* a, const Real* b, size_t N, cudaStream_t stream, pf_delta_stats* out)
{
    // the grid of stats_delta_device (elementwise_blocks, capped): same partials, same order of the final sum, same bits
    const size_t want = (N + CG_THREADS - 1) / CG_THREADS, cap = (size_t) d.sm_count * 16;
    const int blocks = (int) std::min<size_t>(PF_STATIC_STATS_MAX_BLOCKS, std::max<size_t>(1, std::min(want, cap)));
    k_stats_delta<Real><<<blocks, CG_THREADS, 0, stream>>>(N, a, b, d.ptr);
    { cudaError_t e = cudaGetLastError(); if(e != cudaSuccess) return fail(nullptr, PF_ERR_CUDA, "pf_stats_delta_static: launch failed: %s", cudaGetErrorString(e)); }
    pf::StatsPartial h[PF_STATIC_STATS_MAX_BLOCKS];                                       // 40 KB of stack: nothing allocated on the host either
    PF_CUDA(nullptr, cudaMemcpyAsync(h, d.ptr, blocks * sizeof(pf::StatsPartial), cudaMemcpyDeviceToHost, stream));
    PF_CUDA(nullptr, cudaStreamSynchronize(stream));
    Real sum = 0, L1 = 0, L2 = 0, mn = (Real) INFINITY, mx = -(Real) INFINITY;
    for(int i = 0; i < blocks; i++)
    {
        sum += (Real) h[i].sum; L1 += (Real) h[i].L1; L2 += (Real) h[i].L2;
        mn = std::min(mn, (Real) h[i].mn); mx = std::max(mx, (Real) h[i].mx);
    }
    out->sum = (double) sum; out->L1 = (double) L1; out->L2 = (double)(Real) sqrt((double) L2);
    out->min = (double) mn; out->max = (double) mx;
    return PF_OK;
}
} // namespace

extern "C" int pf_stats_delta_static(int device, int dtype, const void* a, const void* b, size_t n, void* stream, pf_delta_stats* out)
{
    static_assert((int) PF_STATIC_STATS_MAX_BLOCKS == (int) CG_MAX_BLOCKS, "same grid cap as pf_stats_delta");
    if(!a || !b || !out) return fail(nullptr, PF_ERR_INVALID, "pf_stats_delta_static: NULL argument");
    if(dtype != PF_F32 && dtype != PF_F64) return fail(nullptr, PF_ERR_INVALID, "pf_stats_delta_static: bad dtype %d", dtype);
    if(device < 0 || device >= STATIC_STATS_MAX_DEVICES) return fail(nullptr, PF_ERR_INVALID, "pf_stats_delta_static: bad device %d", device);
    if(n == 0 || n > (size_t) PF_STATIC_STATS_MAX_ELEMS)
        return fail(nullptr, PF_ERR_INVALID, "pf_stats_delta_static: n = %zu is outside 1 .. PF_STATIC_STATS_MAX_ELEMS = %zu (use pf_stats_delta for larger fields)",
                    n, (size_t) PF_STATIC_STATS_MAX_ELEMS);
    PF_CUDA(nullptr, cudaSetDevice(device));
    StaticStatsDevice d;
    if(int rc = static_stats_setup(device, &d)) return rc;
    return dtype == PF_F64 ? stats_delta_static<double>(d, (const double*) a, (const double*) b, n, (cudaStream_t) stream, out)
                           : stats_delta_static<float>(d, (const float*) a, (const float*) b, n, (cudaStream_t) stream, out);
}
