// pdn_peer_prim.cuh -- peer-memory primitives (flags with epochs, system-scope release/acquire, deadlines) and
// the exchanges as DEVICE functions, so that the kernel that produces a value can also exchange it in its tail:
// an exchange kernel of its own costs 12-20 us per launch (two system fences, a flag round trip and a launch),
// three of them doubled the iteration time on two GPUs.  See pdn_peer.cuh for the protocol.
// Included by pdn_kernels.cuh (namespace pdnk) before the kernels that use it.
#pragma once

constexpr int PEER_MAX = PDN_PEER_MAX;
constexpr size_t PA_HALO_FLAG = PDN_PA_HALO_FLAG, PA_RED_FLAG = PDN_PA_RED_FLAG, PA_RC_FLAG = PDN_PA_RC_FLAG;
constexpr size_t PA_RED_VAL = PDN_PA_RED_VAL, PA_RC_VAL = PDN_PA_RC_VAL, PA_RC_SLOT = PDN_PA_RC_SLOT;
constexpr size_t PA_CW_FLAG = PDN_PA_CW_FLAG, PA_CW_VAL = PDN_PA_CW_VAL;
using PeerLink = PdnPeerLink;

__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// wait until *flag >= want; false (and a status word) after ~2 s
// (the cycles spent waiting are added to *spent: PADNE_B200_PEER_STATS prints them per exchange type)
__device__ __forceinline__ bool peer_wait(const unsigned long long *flag, unsigned long long want, int32_t *status,
                                          unsigned long long *spent) {
    const long long t0 = clock64();
    bool ok = true;
    while (ld_acquire_sys(flag) < want) {
        if (clock64() - t0 > 4000000000LL) {
            atomicExch(&status[0], PDN_ECUDA);
            ok = false;
            break;
        }
    }
    atomicAdd(spent, (unsigned long long)(clock64() - t0));
    return ok;
}

struct HaloSend {
    int32_t off[PEER_MAX], cnt[PEER_MAX], recv_cnt[PEER_MAX];
};

// event log (PADNE_B200_PEER_STATS prints mean intervals between consecutive events of a rank): one thread of
// the exchanging CTA appends (event, globaltimer ns) to a ring behind the epoch words.  Events: 0/1 dots
// exchange enter/exit, 2/3 coarse-residual exchange, 4/5 halo exchange.
constexpr int PEER_LOG = PDN_PEER_LOG;
__device__ __forceinline__ void peer_log(const PeerLink &L, int ev) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    const unsigned long long i = L.epoch[3];
    L.epoch[3] = i + 1;
    L.epoch[8 + (i & (PEER_LOG - 1))] = ((unsigned long long)ev << 56) | (t & ((1ull << 56) - 1));
}

// ---- flag-in-word transport (the "LL" protocol of NCCL) ------------------------------------------------------
// A payload travels as 8-byte words (epoch32 << 32 | data32): an aligned 8-byte store is one transaction, so a
// word whose upper half shows the expected epoch carries valid data -- no system fence after the remote stores, no
// separate flag, no acquire round trip.  Round 1 / early round 2 (stores, __threadfence_system, release flag,
// acquire spin) cost 6 us for three doubles and 9-11 us for the coarse sums on two GPUs; most of it was the fence
// waiting for the NVLink write acknowledgements.  Slots are double-buffered by epoch parity as before.
__device__ __forceinline__ void ll_store(unsigned long long *p, unsigned e32, unsigned v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(((unsigned long long)e32 << 32) | v) : "memory");
}
__device__ __forceinline__ unsigned long long ll_load(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void ll_store64(unsigned long long *p, unsigned e32, unsigned long long v) {
    ll_store(p, e32, (unsigned)v);
    ll_store(p + 1, e32, (unsigned)(v >> 32));
}
// spin until both words of a 64-bit value show epoch e32; false after ~2 s
__device__ __forceinline__ bool ll_wait64(const unsigned long long *p, unsigned e32, unsigned long long &out, int32_t *status,
                                          unsigned long long *spent) {
    unsigned long long a = ll_load(p), b = ll_load(p + 1);
    if ((unsigned)(a >> 32) != e32 || (unsigned)(b >> 32) != e32) {
        const long long t0 = clock64();
        while (true) {
            a = ll_load(p);
            b = ll_load(p + 1);
            if ((unsigned)(a >> 32) == e32 && (unsigned)(b >> 32) == e32) break;
            if (clock64() - t0 > 4000000000LL) { atomicExch(&status[0], PDN_ECUDA); return false; }
        }
        atomicAdd(spent, (unsigned long long)(clock64() - t0));
    }
    out = (a & 0xffffffffull) | (b << 32);
    return true;
}

// ---- all-reduce of three doubles, by ONE WARP (all 32 lanes call, lane 0 holds acc) ----------------------
// Every rank stores its partial sums into a slot on every rank and adds the slots in rank order: identical
// bits everywhere.  Returns false after a deadline miss.  acc comes back valid in lane 0.
// Lanes 0 .. 3 world - 1: lane = 3 k + a stores value a to rank k, then waits for value a of rank k.
__device__ __forceinline__ bool peer_allreduce3(const PeerLink &L, double (&acc)[3]) {
    const int lane = threadIdx.x & 31;
    const unsigned long long e = L.epoch[1] + 1;
    const unsigned e32 = (unsigned)e;
    const size_t slot = (size_t)(e & 1) * PEER_MAX;
    if (lane == 0) peer_log(L, 0);
#pragma unroll
    for (int a = 0; a < 3; a++) acc[a] = __shfl_sync(0xffffffffu, acc[a], 0);
    const int k = lane / 3, a = lane - 3 * k;
    bool ok = true;
    unsigned long long got = 0;
    if (k < L.world) {
        const double mine = a == 0 ? acc[0] : (a == 1 ? acc[1] : acc[2]);
        ll_store64(L.arena[k] + PA_RED_VAL + (slot + L.me) * 8 + 2 * a, e32, (unsigned long long)__double_as_longlong(mine));
        ok = ll_wait64(L.arena[L.me] + PA_RED_VAL + (slot + k) * 8 + 2 * a, e32, got, L.status, L.epoch + 5);
    }
    ok = __all_sync(0xffffffffu, ok);
    // rank order: identical bits on every rank
    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
    for (int r = 0; r < L.world; r++) {
        s0 += __longlong_as_double((long long)__shfl_sync(0xffffffffu, got, 3 * r + 0));
        s1 += __longlong_as_double((long long)__shfl_sync(0xffffffffu, got, 3 * r + 1));
        s2 += __longlong_as_double((long long)__shfl_sync(0xffffffffu, got, 3 * r + 2));
    }
    if (lane == 0) {
        L.epoch[1] = e;
        peer_log(L, 1);
        acc[0] = s0; acc[1] = s1; acc[2] = s2;
    }
    return ok;
}

// ---- exact sum of nc 64-bit integers over the ranks, in place, by ONE CTA (all threads call) --------------
// arange (device, may be null): [first, last] entry this rank can hold a non-zero in -- the aggregates its rows touch
// (agg_range_kernel).  Only that range crosses NVLink (a rank owning 2 of 16 layers: 1/8 of the entries).  Slot
// layout, flag-in-word: words 0, 1 = first, count; then two words per entry.
__device__ __forceinline__ bool peer_sum_i64_cta(const PeerLink &L, long long *rc, int32_t nc, const int32_t *arange) {
    const int tid = threadIdx.x, nt = blockDim.x;
    const unsigned long long e = L.epoch[2] + 1;
    const unsigned e32 = (unsigned)e;
    const size_t slot = (size_t)(e & 1) * PEER_MAX;
    __shared__ int s_lo[PEER_MAX], s_cnt[PEER_MAX], peer_ok;
    if (tid == 0) { peer_log(L, 2); peer_ok = 1; }
    int32_t alo = 0, acnt = nc;
    if (arange != nullptr) {
        const int32_t lo = __ldg(arange), hi = __ldg(arange + 1);
        alo = hi >= lo ? lo : 0;
        acnt = hi >= lo ? hi - lo + 1 : 0;
    }
    if (tid < L.world) {
        unsigned long long *dst = L.arena[tid] + PA_RC_VAL + (slot + L.me) * PA_RC_SLOT;
        ll_store(dst, e32, (unsigned)alo);
        ll_store(dst + 1, e32, (unsigned)acnt);
    }
    for (int j0 = tid; j0 < acnt; j0 += 4 * nt) {      // up to four entries per thread in flight, two stores per peer each
        long long q[4];
#pragma unroll
        for (int u = 0; u < 4; u++) q[u] = j0 + u * nt < acnt ? __ldcg(&rc[alo + j0 + u * nt]) : 0;
        for (int k = 0; k < L.world; k++) {
            unsigned long long *dst = L.arena[k] + PA_RC_VAL + (slot + L.me) * PA_RC_SLOT + 2;
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (j0 + u * nt < acnt) ll_store64(dst + 2 * (size_t)(j0 + u * nt), e32, (unsigned long long)q[u]);
        }
    }
    __syncthreads();      // peer_ok initialised; rc fully read before it is overwritten below
    const unsigned long long *sl = L.arena[L.me] + PA_RC_VAL + slot * PA_RC_SLOT;
    if (tid < L.world) {   // headers of all ranks
        unsigned long long h;
        if (ll_wait64(sl + (size_t)tid * PA_RC_SLOT, e32, h, L.status, L.epoch + 6)) {
            s_lo[tid] = (int)(unsigned)h;
            s_cnt[tid] = (int)(unsigned)(h >> 32);
        } else {
            s_lo[tid] = 0; s_cnt[tid] = 0; peer_ok = 0;
        }
    }
    __syncthreads();
    for (int j = tid; j < nc; j += nt) {
        long long sum = 0;
        bool ok = true;
        for (int r = 0; r < L.world; r++) {       // rank order does not matter for integers; every entry is polled
            const unsigned d = (unsigned)(j - s_lo[r]);
            if (d < (unsigned)s_cnt[r]) {
                unsigned long long v;
                ok = ll_wait64(sl + (size_t)r * PA_RC_SLOT + 2 + 2 * (size_t)d, e32, v, L.status, L.epoch + 6) && ok;
                sum += (long long)v;
            }
        }
        if (!ok) peer_ok = 0;
        rc[j] = sum;
    }
    __syncthreads();
    if (tid == 0) { L.epoch[2] = e; peer_log(L, 3); }
    return peer_ok != 0;
}

// ---- halo: the tail of the kernel that wrote the vector (the LAST CTA, all threads call) ------------------
// every CTA has stored its interface entries into the peers' copies and fenced; raise the flags, then wait
// for the peers' entries of the same epoch
__device__ __forceinline__ void peer_halo_tail(const PeerLink &L, const HaloSend &h) {
    const int tid = threadIdx.x;
    if (tid == 0) peer_log(L, 4);
    __threadfence_system();
    const unsigned long long e = L.epoch[0] + 1;
    if (tid < L.world && tid != L.me && h.cnt[tid] > 0) st_release_sys(L.arena[tid] + PA_HALO_FLAG + L.me, e);
    if (tid < L.world && tid != L.me && h.recv_cnt[tid] > 0)
        peer_wait(L.arena[L.me] + PA_HALO_FLAG + tid, e, L.status, L.epoch + 4);
    __syncthreads();
    if (tid == 0) { L.epoch[0] = e; peer_log(L, 5); }
}

// ---- halo, consumer-side wait (round-2 iteration) ---------------------------------------------------------
// peer_halo_post: the LAST CTA of the kernel that wrote the vector (all threads call) has stored this rank's
// interface entries into the peers' copies: fence, raise the flags, advance the epoch -- and leave.  The wait moves
// to the kernel that READS the vector (peer_halo_wait, first thing after its programmatic-launch wait, threads
// < world of every CTA, followed by a CTA barrier): the flag round trip overlaps the kernel boundary and the
// reader's prologue instead of sitting in the writer's tail (6 us per iteration on 2 GPUs).
__device__ __forceinline__ void peer_halo_post(const PeerLink &L, const HaloSend &h) {
    const int tid = threadIdx.x;
    if (tid == 0) peer_log(L, 4);
    __threadfence_system();
    __syncthreads();
    const unsigned long long e = L.epoch[0] + 1;
    if (tid < L.world && tid != L.me && h.cnt[tid] > 0) st_release_sys(L.arena[tid] + PA_HALO_FLAG + L.me, e);
    __syncthreads();
    if (tid == 0) { L.epoch[0] = e; peer_log(L, 5); }
}
__device__ __forceinline__ void peer_halo_wait(const PeerLink &L, const HaloSend &h) {
    const int tid = threadIdx.x;
    if (tid < L.world && tid != L.me && h.recv_cnt[tid] > 0)
        peer_wait(L.arena[L.me] + PA_HALO_FLAG + tid, L.epoch[0], L.status, L.epoch + 4);
}
