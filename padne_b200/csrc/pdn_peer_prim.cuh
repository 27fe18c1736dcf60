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

// ---- all-reduce of three doubles, by ONE WARP (all 32 lanes call, lane 0 holds acc) ----------------------
// Every rank stores its partial sums into a slot on every rank and adds the slots in rank order: identical
// bits everywhere.  Returns false after a deadline miss.  acc comes back valid in lane 0.
__device__ __forceinline__ bool peer_allreduce3(const PeerLink &L, double (&acc)[3]) {
    const int k = threadIdx.x & 31;
    const unsigned long long e = L.epoch[1] + 1;
    const size_t slot = (size_t)(e & 1) * PEER_MAX;
    if (k == 0) peer_log(L, 0);
#pragma unroll
    for (int a = 0; a < 3; a++) acc[a] = __shfl_sync(0xffffffffu, acc[a], 0);
    if (k < L.world) {
        double *dst = reinterpret_cast<double *>(L.arena[k] + PA_RED_VAL) + (slot + L.me) * 4;
        dst[0] = acc[0]; dst[1] = acc[1]; dst[2] = acc[2];
        __threadfence_system();
        st_release_sys(L.arena[k] + PA_RED_FLAG + slot + L.me, e);
    }
    bool ok = true;
    if (k < L.world) ok = peer_wait(L.arena[L.me] + PA_RED_FLAG + slot + k, e, L.status, L.epoch + 5);
    ok = __all_sync(0xffffffffu, ok);
    if (k == 0) {
        L.epoch[1] = e;
        peer_log(L, 1);
        acc[0] = acc[1] = acc[2] = 0.0;
        const volatile double *src = reinterpret_cast<const volatile double *>(L.arena[L.me] + PA_RED_VAL) + slot * 4;
        for (int r = 0; r < L.world; r++) {
            acc[0] += src[r * 4 + 0]; acc[1] += src[r * 4 + 1]; acc[2] += src[r * 4 + 2];
        }
    }
    return ok;
}

// ---- exact sum of nc 64-bit integers over the ranks, in place, by ONE CTA (all threads call) --------------
__device__ __forceinline__ bool peer_sum_i64_cta(const PeerLink &L, long long *rc, int32_t nc) {
    const int tid = threadIdx.x, nt = blockDim.x;
    const unsigned long long e = L.epoch[2] + 1;
    const size_t slot = (size_t)(e & 1) * PEER_MAX;
    if (tid == 0) peer_log(L, 2);
    for (int k = 0; k < L.world; k++) {
        long long *dst = reinterpret_cast<long long *>(L.arena[k] + PA_RC_VAL) + (slot + L.me) * PA_RC_SLOT;
        for (int j = tid; j < nc; j += nt) dst[j] = __ldcg(&rc[j]);
    }
    __threadfence_system();
    __syncthreads();
    if (tid < L.world) st_release_sys(L.arena[tid] + PA_RC_FLAG + slot + L.me, e);
    __shared__ int peer_ok;
    if (tid == 0) peer_ok = 1;
    __syncthreads();
    if (tid < L.world && !peer_wait(L.arena[L.me] + PA_RC_FLAG + slot + tid, e, L.status, L.epoch + 6)) peer_ok = 0;
    __syncthreads();
    if (tid == 0) L.epoch[2] = e;
    if (!peer_ok) return false;
    const volatile long long *src = reinterpret_cast<const volatile long long *>(L.arena[L.me] + PA_RC_VAL) + slot * PA_RC_SLOT;
    for (int j = tid; j < nc; j += nt) {
        long long sum = 0;
        for (int r = 0; r < L.world; r++) sum += src[(size_t)r * PA_RC_SLOT + j];
        rc[j] = sum;
    }
    if (tid == 0) peer_log(L, 3);
    return true;
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
