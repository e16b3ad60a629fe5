// One-shot exchange over NVLink peer memory -- the step's only cross-rank traffic.
//
// Every rank owns one buffer of tac_xchg_bytes() in peer-mappable memory, mapped in every process of the node:
//
//   [0, 256)    control words (local use only): seq_push, seq_pop, push_done, pop_done, wait_error
//   [256, 512)  flags [2 parities][8 source ranks]  uint32, written by the SOURCE rank
//   [512, ...)  slots [2 parities][world][payload]  float64, slot r written by rank r
//
// push: a rank peer-stores its payload into slot[parity][rank] of EVERY rank (its own included), fences at system
// scope and then stores flag[parity][rank] = seq + 1 on every rank.  pop: a rank spins on its own world flags and adds
// the world slots in RANK ORDER -- every rank forms the bit-identical float64 sum, with no reduction kernel of a
// communication library to schedule and no SM to reserve for it.  parity = seq & 1: a rank can be at most one push
// ahead of the slowest rank's pop (it needs that rank's next push to finish its own pop), so two slots suffice.
// The sequence counters live on the device and advance with every execution, eager or CUDA-graph replay alike.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tac_spectral.h"

namespace tac {

constexpr int kXchgMaxWorld = TAC_XCHG_MAX_WORLD;
constexpr size_t kXchgFlagOff = 256, kXchgSlotOff = 512;

struct XchgDev {                       // by-value kernel argument; world <= 1: no exchange
    int world, rank;
    long long payload;                 // float64 per slot
    unsigned long long timeout_ns;     // bound of one xchg_wait (0: wait for ever), see xchg_timeout_ns()
    unsigned char* peer[kXchgMaxWorld];
};

// How long a pop waits for a peer's push before it gives up: tac_xchg_set_timeout(), else TAC_XCHG_TIMEOUT_S, else
// 600 s (ranks of a training job drift apart by whole seconds to minutes: checkpoints, validation, a stalled loader).
unsigned long long xchg_timeout_ns();   // exchange.cu

inline XchgDev xchg_dev(const tac_xchg* x) {
    XchgDev d;
    d.world = 1; d.rank = 0; d.payload = 0; d.timeout_ns = 0;
    for (int i = 0; i < kXchgMaxWorld; ++i) d.peer[i] = nullptr;
    if (x && x->world > 1) {
        d.world = x->world; d.rank = x->rank; d.payload = x->payload;
        d.timeout_ns = xchg_timeout_ns();
        for (int i = 0; i < x->world; ++i) d.peer[i] = reinterpret_cast<unsigned char*>(x->peer[i]);
    }
    return d;
}

__device__ __forceinline__ unsigned* xchg_ctl(const XchgDev& x) { return reinterpret_cast<unsigned*>(x.peer[x.rank]); }
__device__ __forceinline__ unsigned* xchg_flags(unsigned char* base, int parity) {
    return reinterpret_cast<unsigned*>(base + kXchgFlagOff) + parity * kXchgMaxWorld;
}
__device__ __forceinline__ double* xchg_slot(unsigned char* base, const XchgDev& x, int parity, int src) {
    return reinterpret_cast<double*>(base + kXchgSlotOff) + ((size_t)parity * x.world + src) * (size_t)x.payload;
}

__device__ __forceinline__ unsigned xchg_ld_volatile(const unsigned* p) {
    unsigned v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned xchg_ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void xchg_st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

// sequence number of the push this kernel performs (read by every CTA before it signs off in xchg_push_commit)
__device__ __forceinline__ unsigned xchg_push_seq(const XchgDev& x) { return xchg_ld_volatile(xchg_ctl(x) + 0); }
__device__ __forceinline__ unsigned xchg_pop_seq(const XchgDev& x) { return xchg_ld_volatile(xchg_ctl(x) + 1); }

// value idx of this rank's payload -> slot[parity][rank] on every rank
__device__ __forceinline__ void xchg_store(const XchgDev& x, unsigned seq, size_t idx, double v) {
    const int parity = (int)(seq & 1u);
#pragma unroll 1
    for (int p = 0; p < x.world; ++p) xchg_slot(x.peer[p], x, parity, x.rank)[idx] = v;
}

// Called by ALL threads of every CTA of the pushing kernel, after their xchg_store()s.  The last CTA raises the flags.
__device__ __forceinline__ void xchg_push_commit(const XchgDev& x, unsigned seq, unsigned nctas) {
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned* ctl = xchg_ctl(x);
        const unsigned prev = atomicAdd(ctl + 2, 1u);
        if (prev == nctas - 1) {
            ctl[2] = 0;
            __threadfence_system();
            const int parity = (int)(seq & 1u);
#pragma unroll 1
            for (int p = 0; p < x.world; ++p) xchg_st_release_sys(xchg_flags(x.peer[p], parity) + x.rank, seq + 1u);
            ctl[0] = seq + 1u;
        }
    }
}

__device__ __forceinline__ unsigned long long xchg_globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Threads 0 .. world-1 spin on the local flags of pop number `seq`; the caller follows with __syncthreads().
// Bounded by x.timeout_ns of wall-clock time (0: unbounded).  A wait that runs out does NOT trap -- a trap is a sticky
// context error that takes the rank down, and its peers after it: it records `seq + 1` in control word 4 of the rank's
// own buffer (tac_xchg_status reads it on the host) and lets the kernel finish on whatever the slots hold.  From then on
// the exchange's results are undefined; the host is expected to treat a non-zero status as fatal for the job.
__device__ __forceinline__ void xchg_wait(const XchgDev& x, unsigned seq) {
    if ((int)threadIdx.x < x.world) {
        const unsigned* f = xchg_flags(x.peer[x.rank], (int)(seq & 1u)) + threadIdx.x;
        const unsigned long long t0 = xchg_globaltimer_ns();
        unsigned spins = 0;
        while ((int)(xchg_ld_acquire_sys(f) - (seq + 1u)) < 0) {
            if (x.timeout_ns && (++spins & 63u) == 0 && xchg_globaltimer_ns() - t0 > x.timeout_ns) {
                atomicMax(xchg_ctl(x) + 4, seq + 1u);
                break;
            }
            __nanosleep(64);
        }
    }
}

// slot reads go to L2 (the peers' stores land there; this SM's L1 may still hold the slot's contents of two pops ago)
__device__ __forceinline__ double xchg_sum(const XchgDev& x, unsigned seq, size_t idx) {
    const int parity = (int)(seq & 1u);
    double acc = 0.0;
#pragma unroll 1
    for (int r = 0; r < x.world; ++r) acc += __ldcg(xchg_slot(x.peer[x.rank], x, parity, r) + idx);   // fixed rank order
    return acc;
}

// Called by thread 0 of every CTA of the popping kernel once the CTA has read its slots.
__device__ __forceinline__ void xchg_pop_commit(const XchgDev& x, unsigned seq, unsigned nctas) {
    unsigned* ctl = xchg_ctl(x);
    const unsigned prev = atomicAdd(ctl + 3, 1u);
    if (prev == nctas - 1) { ctl[3] = 0; __threadfence(); ctl[1] = seq + 1u; }
}

}  // namespace tac
