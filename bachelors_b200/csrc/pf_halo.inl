// pf_halo.inl - host side of the device-initiated halo exchange (pf_halo.cuh). Included at the end of pf_solver.cu.

extern "C" {

int pf_ipc_export(pf_solver* s, void* dev_ptr, void* handle)
{
    if(int rc = bind(s)) return rc;
    if(!dev_ptr || !handle) return fail(s, PF_ERR_INVALID, "pf_ipc_export: NULL argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == PF_IPC_HANDLE_BYTES, "cudaIpcMemHandle_t size");
    cudaIpcMemHandle_t h;
    PF_CUDA(s, cudaIpcGetMemHandle(&h, dev_ptr));
    memcpy(handle, &h, sizeof h);
    return PF_OK;
}

int pf_ipc_open(pf_solver* s, const void* handle, void** mapped)
{
    if(int rc = bind(s)) return rc;
    if(!handle || !mapped) return fail(s, PF_ERR_INVALID, "pf_ipc_open: NULL argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    *mapped = nullptr;
    PF_CUDA(s, cudaIpcOpenMemHandle(mapped, h, cudaIpcMemLazyEnablePeerAccess));
    return PF_OK;
}

int pf_ipc_close(pf_solver* s, void* mapped)
{
    if(int rc = bind(s)) return rc;
    if(mapped) PF_CUDA(s, cudaIpcCloseMemHandle(mapped));
    return PF_OK;
}

void* pf_halo_flags(pf_solver* s)
{
    if(!s || bind(s)) return nullptr;
    if(!s->halo_flags)
    {
        if(cudaMalloc((void**) &s->halo_flags, HALO_FLAG_WORDS * sizeof(unsigned)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        cudaMemset(s->halo_flags, 0, HALO_FLAG_WORDS * sizeof(unsigned));
    }
    return s->halo_flags;
}

int pf_halo_push_async(pf_solver* s, int side, int n_fields, const void* const* fields, void* const* peer_fields, size_t peer_first_row,
                       void* peer_flags, int peer_flag_index, unsigned value, int on_aux_stream)
{
    if(int rc = bind(s)) return rc;
    if(n_fields < 1 || n_fields > 4 || !fields || !peer_fields || !peer_flags || (side != 0 && side != 1) ||
       peer_flag_index < 0 || peer_flag_index > HALO_FLAG_FROM_HI)
        return fail(s, PF_ERR_INVALID, "pf_halo_push_async: bad arguments");
    if(!(side == 0 ? s->slab.has_lo : s->slab.has_hi)) return PF_OK;
    if(!pf_halo_flags(s)) return fail(s, PF_ERR_OOM, "pf_halo_push_async: no flag block");
    const size_t rs = s->real_size(), row_bytes = (size_t) s->nx * rs;
    const int g = s->slab.ghost;
    HaloPushArgs a; memset(&a, 0, sizeof a);
    a.n_fields = n_fields;
    a.bytes = (size_t) g * row_bytes;
    if(a.bytes % 16) return fail(s, PF_ERR_UNSUPPORTED, "pf_halo_push_async: ghost rows of %zu bytes are not a multiple of 16", a.bytes);
    const size_t src_row = side == 0 ? (size_t) s->ghost_lo : (size_t)(s->ghost_lo + s->slab.ny_local - g);
    for(int i = 0; i < n_fields; i++)
    {
        if(!fields[i] || !peer_fields[i]) return fail(s, PF_ERR_INVALID, "pf_halo_push_async: NULL field");
        a.src[i] = (const char*) fields[i] + src_row * row_bytes;
        a.dst[i] = (char*) peer_fields[i] + peer_first_row * row_bytes;
    }
    a.peer_flag = (unsigned*) peer_flags + peer_flag_index;
    a.value = value;
    a.ticket = s->halo_flags + HALO_FLAG_TICKET + side;                   // one counter per side: the two pushes of a step may overlap
    const int blocks = (int) std::max<size_t>(1, std::min<size_t>((a.bytes / 16 + 255) / 256, (size_t) s->sm_count));
    k_halo_push<<<blocks, 256, 0, on_aux_stream ? s->aux_stream : s->stream>>>(a);
    PF_LAUNCH_CHECK(s);
    return PF_OK;
}

int pf_halo_wait_async(pf_solver* s, int flag_index, unsigned value, int on_aux_stream)
{
    if(int rc = bind(s)) return rc;
    if(flag_index < 0 || flag_index > HALO_FLAG_FROM_HI) return fail(s, PF_ERR_INVALID, "pf_halo_wait_async: flag %d", flag_index);
    if(!pf_halo_flags(s)) return fail(s, PF_ERR_OOM, "pf_halo_wait_async: no flag block");
    k_halo_wait<<<1, 1, 0, on_aux_stream ? s->aux_stream : s->stream>>>(s->halo_flags, flag_index, value);
    PF_LAUNCH_CHECK(s);
    return PF_OK;
}

int pf_halo_link(pf_solver* s, int n_fields, void* const* fields, void* const* lo_fields, size_t lo_first_row, void* lo_flags,
                 void* const* hi_fields, void* hi_flags)
{
    if(int rc = bind(s)) return rc;
    s->halo.on = false;
    if(n_fields == 0) return PF_OK;                                        // unlink
    if(n_fields < 2 || n_fields > 8 || !fields) return fail(s, PF_ERR_INVALID, "pf_halo_link: bad field list");
    if((lo_flags != nullptr) != (s->slab.has_lo != 0) || (hi_flags != nullptr) != (s->slab.has_hi != 0))
        return fail(s, PF_ERR_INVALID, "pf_halo_link: a neighbour must be linked exactly where the slab has one");
    if((lo_flags && !lo_fields) || (hi_flags && !hi_fields)) return fail(s, PF_ERR_INVALID, "pf_halo_link: NULL peer field list");
    if(!pf_halo_flags(s)) return fail(s, PF_ERR_OOM, "pf_halo_link: no flag block");
    if(((size_t) s->slab.ghost * (size_t) s->nx * s->real_size()) % 16) return fail(s, PF_ERR_UNSUPPORTED, "pf_halo_link: ghost rows are not a multiple of 16 bytes");
    s->halo.n_fields = n_fields;
    for(int i = 0; i < n_fields; i++)
    {
        s->halo.local[i] = fields[i];
        s->halo.peer[0][i] = lo_flags ? lo_fields[i] : nullptr;
        s->halo.peer[1][i] = hi_flags ? hi_fields[i] : nullptr;
    }
    s->halo.first_row[0] = lo_first_row; s->halo.first_row[1] = 0;
    s->halo.peer_flags[0] = (unsigned*) lo_flags; s->halo.peer_flags[1] = (unsigned*) hi_flags;
    s->halo.e = 0;
    PF_CUDA(s, cudaMemset(s->halo_flags, 0, HALO_FLAG_WORDS * sizeof(unsigned)));
    s->halo.on = true;
    return PF_OK;
}

int pf_halo_check(pf_solver* s)
{
    if(int rc = bind(s)) return rc;
    if(!s->halo_flags) return PF_OK;
    unsigned stalled = 0;
    PF_CUDA(s, cudaMemcpy(&stalled, s->halo_flags + HALO_FLAG_STALLED, sizeof stalled, cudaMemcpyDeviceToHost));
    if(stalled) return fail(s, PF_ERR_CUDA, "halo exchange: waited in vain for the %s neighbour's rows", stalled == 1 ? "lower" : "upper");
    return PF_OK;
}

} // extern "C"
