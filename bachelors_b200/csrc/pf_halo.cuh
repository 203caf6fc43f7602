// pf_halo.cuh - device-initiated halo exchange between the slabs of one box over NVLink peer memory.
//
// The reference is single-GPU; the first version of the slab layer handed the edge rows to NCCL send/recv. NCCL's kernels need
// SM resources that the interior kernel - one wave of long-lived CTAs that fill the register file - does not release until it
// ends, so the "overlapped" exchange really ran after the interior (measured: 43 us of a 518 us step at 2 GPUs, 75 us at 8).
// Here a rank stores its freshly computed edge rows straight into the neighbour's ghost rows through a peer mapping
// (cudaIpc) with a small copy kernel that fits beside the interior CTAs, then publishes the exchange number into a flag word
// in the neighbour's memory; the neighbour's next edge-row launch is preceded by a one-thread kernel that acquires that flag.
// No NCCL call, no host synchronisation, no Python between the ranks.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pf {

enum { HALO_FLAG_FROM_LO = 0, HALO_FLAG_FROM_HI = 1, HALO_FLAG_STALLED = 2, HALO_FLAG_TICKET = 3, HALO_FLAG_WORDS = 8 };

struct HaloPushArgs {
    const void* src[4]; void* dst[4];      // per field: first source element / first destination element (peer memory)
    int n_fields;
    size_t bytes;                          // per field, a multiple of 16
    unsigned* peer_flag;                   // word in the PEER's flag block
    unsigned value;
    unsigned* ticket;                      // local word: arrival counter of this launch's blocks
};

// Copies n_fields x bytes with 16-byte accesses; the last block to finish publishes `value` (system-scope release).
__global__ void __launch_bounds__(256)
k_halo_push(const HaloPushArgs a)
{
    const size_t n16 = a.bytes / 16, stride = (size_t) gridDim.x * blockDim.x;
    for(int f = 0; f < a.n_fields; f++)
    {
        const uint4* s = reinterpret_cast<const uint4*>(a.src[f]);
        uint4* d = reinterpret_cast<uint4*>(a.dst[f]);
        for(size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n16; i += stride) d[i] = s[i];
    }
    __syncthreads();
    if(threadIdx.x == 0)
    {
        __threadfence_system();                                             // this block's peer stores before its ticket
        if(atomicAdd(a.ticket, 1u) == gridDim.x - 1)
        {
            *a.ticket = 0;
            __threadfence_system();
            asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(a.peer_flag), "r"(a.value) : "memory");
        }
    }
}

// One thread: waits until the neighbour has published exchange number `value` (or later) into this rank's flag word.
__global__ void k_halo_wait(unsigned* flags, int which, unsigned value)
{
    unsigned v, spins = 0;
    for(;;)
    {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + which) : "memory");
        if((int)(v - value) >= 0) break;
        __nanosleep(100);
        if(++spins > (1u << 24)) { flags[HALO_FLAG_STALLED] = 1u + (unsigned) which; break; }     // ~ 2 s: the neighbour is gone; the host reports it
    }
}

} // namespace pf
