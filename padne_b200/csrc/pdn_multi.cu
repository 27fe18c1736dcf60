// pdn_multi.cu -- multi-GPU plumbing: one process per GPU, NCCL over NVLink / NVSwitch.
//
// The north star partitions one large board across the GPUs of a box: contiguous row ranges of the
// (permuted) reduced system, a halo exchange of the interface x-values before every SpMV and an
// all-reduce of the fused Krylov dot products.  This file owns the communicator and the three
// collective shapes the solver needs; the Krylov drivers in pdn_solve.cu enqueue them on the context
// stream (they are captured into the same CUDA graph as the kernels).
//
// NCCL is resolved at run time with dlopen("libnccl.so.2"): inside a torch process that is the copy
// torch already loaded (2.28.x), otherwise the system library.  The few NCCL types used here are
// re-declared so the build needs no NCCL headers and the library no link-time NCCL dependency.
#include <dlfcn.h>
#include <stdlib.h>

#include "pdn_common.cuh"

namespace {

// --- the slice of the NCCL ABI used here (nccl.h 2.27: enums and opaque types are stable) ----------
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
constexpr int kNcclSuccess = 0;
constexpr int kNcclSum = 0;
constexpr int kNcclMax = 2;       // ncclMax
constexpr int kNcclInt32 = 2;     // ncclInt32 / ncclInt
constexpr int kNcclInt8 = 0;      // ncclInt8 / ncclChar
constexpr int kNcclInt64 = 4;     // ncclInt64
constexpr int kNcclFloat64 = 8;   // ncclFloat64

struct NcclApi {
    void *handle = nullptr;
    int (*GetUniqueId)(ncclUniqueId *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Broadcast)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Send)(const void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};

NcclApi g_nccl;

const NcclApi &nccl() {
    if (g_nccl.handle) return g_nccl;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) PDN_FAIL(PDN_ECUDA, "comm: cannot load libnccl.so.2 (%s)", dlerror());
#define PDN_SYM(name)                                                                    \
    do {                                                                                 \
        *(void **)(&g_nccl.name) = dlsym(h, "nccl" #name);                               \
        if (!g_nccl.name) PDN_FAIL(PDN_ECUDA, "comm: libnccl lacks nccl" #name);          \
    } while (0)
    PDN_SYM(GetUniqueId);
    PDN_SYM(CommInitRank);
    PDN_SYM(CommDestroy);
    PDN_SYM(AllReduce);
    PDN_SYM(Broadcast);
    PDN_SYM(AllGather);
    PDN_SYM(Send);
    PDN_SYM(Recv);
    PDN_SYM(GroupStart);
    PDN_SYM(GroupEnd);
    PDN_SYM(GetErrorString);
#undef PDN_SYM
    g_nccl.handle = h;
    return g_nccl;
}

#define PDN_NCCL(expr)                                                                            \
    do {                                                                                          \
        int _r = (expr);                                                                          \
        if (_r != kNcclSuccess)                                                                   \
            PDN_FAIL(PDN_ECUDA, "%s:%d: %s -> %s", __FILE__, __LINE__, #expr, nccl().GetErrorString(_r)); \
    } while (0)

}  // namespace

struct PdnComm {
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    // peer-memory exchanges (pdn_peer.cuh): one arena per rank, mapped into every peer through CUDA IPC
    bool p2p = false;
    unsigned long long *arena = nullptr;              // local
    unsigned long long *peer[PDN_PEER_MAX] = {};      // peer[k] = rank k's arena mapped here (peer[rank] = arena)
    int64_t n_cap = 0;                                // doubles per exchanged vector
    unsigned long long *epoch = nullptr;              // device [8 + log]
    PdnPeerLink *link_d = nullptr;                    // device copy of pdn_peer_link() (kernels take a pointer)
    unsigned int *ticket = nullptr;                   // device [4]
};

static void peer_unmap(PdnComm *c) {
    for (int k = 0; k < c->world && k < PDN_PEER_MAX; k++) {
        if (k != c->rank && c->peer[k]) cudaIpcCloseMemHandle(c->peer[k]);
        c->peer[k] = nullptr;
    }
    if (c->arena) cudaFree(c->arena);
    c->arena = nullptr;
    c->n_cap = 0;
}

bool pdn_peer_enabled(const pdn_ctx *ctx) { return ctx->comm && ctx->comm->p2p; }

void pdn_peer_ensure(pdn_ctx *ctx, int64_t n) {
    PdnComm *c = ctx->comm;
    if (!c || !c->p2p || n <= c->n_cap) return;
    cudaStream_t st = ctx->stream;
    PDN_CUDA(cudaStreamSynchronize(st));
    peer_unmap(c);
    const int64_t n_cap = (n + n / 8 + 1024 + 31) / 32 * 32;
    const size_t words = PDN_PA_VEC + (size_t)PDN_PA_NVEC * n_cap;
    PDN_CUDA(cudaMalloc((void **)&c->arena, words * 8));
    PDN_CUDA(cudaMemsetAsync(c->arena, 0, words * 8, st));
    PDN_CUDA(cudaMemsetAsync(c->epoch, 0, 4 * sizeof(unsigned long long), st));   // the wait-cycle counters [4..6] keep counting
    PDN_CUDA(cudaMemsetAsync(c->epoch + 7, 0, sizeof(unsigned long long), st));   // epoch of the merged dots + c_w exchange
    PDN_CUDA(cudaMemsetAsync(c->ticket, 0, 4 * sizeof(unsigned int), st));
    // handles of all ranks (the all-gather also orders every rank's memset before any peer store)
    cudaIpcMemHandle_t mine;
    PDN_CUDA(cudaIpcGetMemHandle(&mine, c->arena));
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    DevBuf<char> hbuf;
    hbuf.reserve((size_t)64 * c->world);
    PDN_CUDA(cudaMemcpyAsync(hbuf.p + 64 * c->rank, &mine, 64, cudaMemcpyHostToDevice, st));
    PDN_NCCL(nccl().AllGather(hbuf.p + 64 * c->rank, hbuf.p, 64, kNcclInt8, c->comm, st));
    std::vector<cudaIpcMemHandle_t> all(c->world);
    PDN_CUDA(cudaMemcpyAsync(all.data(), hbuf.p, (size_t)64 * c->world, cudaMemcpyDeviceToHost, st));
    PDN_CUDA(cudaStreamSynchronize(st));
    for (int k = 0; k < c->world; k++) {
        if (k == c->rank) { c->peer[k] = c->arena; continue; }
        void *ptr = nullptr;
        PDN_CUDA(cudaIpcOpenMemHandle(&ptr, all[k], cudaIpcMemLazyEnablePeerAccess));
        c->peer[k] = (unsigned long long *)ptr;
    }
    c->n_cap = n_cap;
    if (!c->link_d) PDN_CUDA(cudaMalloc((void **)&c->link_d, sizeof(PdnPeerLink)));
    const PdnPeerLink L = pdn_peer_link(ctx);
    PDN_CUDA(cudaMemcpy(c->link_d, &L, sizeof L, cudaMemcpyHostToDevice));
}
const PdnPeerLink *pdn_peer_link_dev(const pdn_ctx *ctx) { return ctx->comm ? ctx->comm->link_d : nullptr; }

PdnPeerLink pdn_peer_link(const pdn_ctx *ctx) {
    PdnPeerLink L{};
    const PdnComm *c = ctx->comm;
    L.world = c->world;
    L.me = c->rank;
    for (int k = 0; k < c->world; k++) L.arena[k] = c->peer[k];
    L.epoch = c->epoch;
    L.status = ctx->d_status;
    return L;
}
size_t pdn_peer_vec_offset(const pdn_ctx *ctx, int which) { return PDN_PA_VEC + (size_t)which * ctx->comm->n_cap; }
double *pdn_peer_vec(const pdn_ctx *ctx, int which) {
    return reinterpret_cast<double *>(ctx->comm->arena + pdn_peer_vec_offset(ctx, which));
}
unsigned int *pdn_peer_ticket(const pdn_ctx *ctx) { return ctx->comm->ticket; }

int pdn_comm_rank(const pdn_ctx *ctx) { return ctx->comm ? ctx->comm->rank : 0; }
int pdn_comm_world(const pdn_ctx *ctx) { return ctx->comm ? ctx->comm->world : 1; }

void pdn_comm_release(pdn_ctx *ctx) {
    if (ctx->comm) {
        if (ctx->comm->epoch && getenv("PADNE_B200_PEER_STATS")) {
            unsigned long long h[8] = {};
            cudaMemcpy(h, ctx->comm->epoch, sizeof h, cudaMemcpyDeviceToHost);
            fprintf(stderr, "[padne_b200 rank %d] exchanges: halo %llu, dots %llu, coarse %llu; cycles waited: halo %llu, dots %llu, coarse %llu\n",
                    ctx->comm->rank, h[0], h[1], h[2], h[4], h[5], h[6]);
            // event log: mean time between consecutive events (0/1 dots enter/exit, 2/3 coarse, 4/5 halo)
            std::vector<unsigned long long> lg(PDN_PEER_LOG);
            cudaMemcpy(lg.data(), ctx->comm->epoch + 8, lg.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
            const unsigned long long cnt = h[3], nlog = cnt < (unsigned long long)PDN_PEER_LOG ? cnt : PDN_PEER_LOG;
            double sum[8][8] = {};
            int num[8][8] = {};
            for (unsigned long long q = cnt - nlog + 1; q < cnt; q++) {
                const unsigned long long a = lg[(q - 1) & (PDN_PEER_LOG - 1)], b = lg[q & (PDN_PEER_LOG - 1)];
                const int ea = (int)(a >> 56) & 7, eb = (int)(b >> 56) & 7;
                const double dt = (double)((b & ((1ull << 56) - 1)) - (a & ((1ull << 56) - 1))) * 1e-3;
                if (dt < 0 || dt > 1e4) continue;   // between solves
                sum[ea][eb] += dt;
                num[ea][eb]++;
            }
            std::string out;
            char buf[64];
            for (int a = 0; a < 8; a++)
                for (int b = 0; b < 8; b++)
                    if (num[a][b] > 16) {
                        snprintf(buf, sizeof buf, " %d->%d %.1f us (n=%d)", a, b, sum[a][b] / num[a][b], num[a][b]);
                        out += buf;
                    }
            fprintf(stderr, "[padne_b200 rank %d] event intervals:%s\n", ctx->comm->rank, out.c_str());
        }
        peer_unmap(ctx->comm);
        if (ctx->comm->epoch) cudaFree(ctx->comm->epoch);
        if (ctx->comm->link_d) cudaFree(ctx->comm->link_d);
        if (ctx->comm->ticket) cudaFree(ctx->comm->ticket);
        if (ctx->comm->comm) nccl().CommDestroy(ctx->comm->comm);
        delete ctx->comm;
        ctx->comm = nullptr;
    }
}

// in-place sum of `count` doubles on the context stream
void pdn_comm_allreduce_sum(pdn_ctx *ctx, double *buf, int count) {
    PDN_NCCL(nccl().AllReduce(buf, buf, (size_t)count, kNcclFloat64, kNcclSum, ctx->comm->comm, ctx->stream));
}

// in-place sum of 64-bit integers (the fixed-point coarse residual: exact, order-independent)
void pdn_comm_allreduce_sum_i64(pdn_ctx *ctx, long long *buf, int count) {
    PDN_NCCL(nccl().AllReduce(buf, buf, (size_t)count, kNcclInt64, kNcclSum, ctx->comm->comm, ctx->stream));
}

// in-place maximum of `count` doubles over the ranks
void pdn_comm_allreduce_max(pdn_ctx *ctx, double *buf, int count) {
    PDN_NCCL(nccl().AllReduce(buf, buf, (size_t)count, kNcclFloat64, kNcclMax, ctx->comm->comm, ctx->stream));
}

// in-place sum of 32-bit integers (halo marks: every rank sets its own bit, so the sum is the bitwise OR)
void pdn_comm_allreduce_sum_i32(pdn_ctx *ctx, int32_t *buf, size_t count) {
    PDN_NCCL(nccl().AllReduce(buf, buf, count, kNcclInt32, kNcclSum, ctx->comm->comm, ctx->stream));
}

// int32 vec[begin[k] .. begin[k+1]) is broadcast from rank k, for every k
void pdn_comm_allgatherv_i32(pdn_ctx *ctx, int32_t *vec, const int64_t *begin) {
    const int world = ctx->comm->world;
    PDN_NCCL(nccl().GroupStart());
    for (int k = 0; k < world; k++) {
        const int64_t cnt = begin[k + 1] - begin[k];
        if (cnt > 0)
            PDN_NCCL(nccl().Broadcast(vec + begin[k], vec + begin[k], (size_t)cnt, kNcclInt32, k, ctx->comm->comm, ctx->stream));
    }
    PDN_NCCL(nccl().GroupEnd());
}

// vec[begin[k] .. begin[k+1]) is broadcast from rank k, for every k (an all-gather with ragged slices)
void pdn_comm_allgatherv(pdn_ctx *ctx, double *vec, const int64_t *begin) {
    const int world = ctx->comm->world;
    PDN_NCCL(nccl().GroupStart());
    for (int k = 0; k < world; k++) {
        const int64_t cnt = begin[k + 1] - begin[k];
        if (cnt > 0)
            PDN_NCCL(nccl().Broadcast(vec + begin[k], vec + begin[k], (size_t)cnt, kNcclFloat64, k, ctx->comm->comm,
                                      ctx->stream));
    }
    PDN_NCCL(nccl().GroupEnd());
}

// grouped point-to-point exchange: to/from peer k, send_cnt[k] doubles at sendbuf + send_off[k] and
// recv_cnt[k] doubles at recvbuf + recv_off[k]
void pdn_comm_exchange(pdn_ctx *ctx, const double *sendbuf, const int32_t *send_off, const int32_t *send_cnt,
                       double *recvbuf, const int32_t *recv_off, const int32_t *recv_cnt) {
    const int world = ctx->comm->world, me = ctx->comm->rank;
    PDN_NCCL(nccl().GroupStart());
    for (int k = 0; k < world; k++) {
        if (k == me) continue;
        if (send_cnt[k] > 0)
            PDN_NCCL(nccl().Send(sendbuf + send_off[k], (size_t)send_cnt[k], kNcclFloat64, k, ctx->comm->comm, ctx->stream));
        if (recv_cnt[k] > 0)
            PDN_NCCL(nccl().Recv(recvbuf + recv_off[k], (size_t)recv_cnt[k], kNcclFloat64, k, ctx->comm->comm, ctx->stream));
    }
    PDN_NCCL(nccl().GroupEnd());
}

extern "C" {

int pdn_comm_unique_id(void *out128) {
    if (!out128) return PDN_EINVAL;
    try {
        ncclUniqueId id;
        PDN_NCCL(nccl().GetUniqueId(&id));
        memcpy(out128, &id, sizeof id);
        return PDN_OK;
    } catch (const pdn_error &e) {
        fprintf(stderr, "pdn_comm_unique_id: %s\n", e.msg.c_str());
        return e.code;
    }
}

int pdn_comm_init(pdn_ctx *ctx, int rank, int world, const void *id128) {
    return pdn_guard(ctx, [&] {
        if (world < 1 || world > 32 || rank < 0 || rank >= world || !id128)
            PDN_FAIL(PDN_EINVAL, "comm_init: rank %d of %d (1..32 ranks supported)", rank, world);
        pdn_comm_release(ctx);
        if (world == 1) return;
        ncclUniqueId id;
        memcpy(&id, id128, sizeof id);
        PdnComm *c = new PdnComm();
        c->rank = rank;
        c->world = world;
        ctx->comm = c;
        PDN_NCCL(nccl().CommInitRank(&c->comm, world, id, rank));
        // first use sets up the channels; do it outside any graph capture
        DevBuf<double> tmp;
        tmp.reserve(8);
        PDN_CUDA(cudaMemsetAsync(tmp.p, 0, 8 * sizeof(double), ctx->stream));
        pdn_comm_allreduce_sum(ctx, tmp.p, 4);
        PDN_CUDA(cudaStreamSynchronize(ctx->stream));
        // peer-memory exchanges unless PADNE_B200_COMM=nccl (or more ranks than the arena has slots)
        const char *mode = getenv("PADNE_B200_COMM");
        c->p2p = world <= PDN_PEER_MAX && !(mode && strcmp(mode, "nccl") == 0);
        if (c->p2p) {
            PDN_CUDA(cudaMalloc((void **)&c->epoch, (8 + PDN_PEER_LOG) * sizeof(unsigned long long)));
            PDN_CUDA(cudaMemset(c->epoch, 0, (8 + PDN_PEER_LOG) * sizeof(unsigned long long)));
            PDN_CUDA(cudaMalloc((void **)&c->ticket, 4 * sizeof(unsigned int)));
        }
    });
}

int pdn_comm_destroy(pdn_ctx *ctx) {
    return pdn_guard(ctx, [&] {
        PDN_CUDA(cudaStreamSynchronize(ctx->stream));
        pdn_comm_release(ctx);
    });
}

int pdn_comm_info(pdn_ctx *ctx, int *rank, int *world) {
    if (!ctx) return PDN_EINVAL;
    if (rank) *rank = pdn_comm_rank(ctx);
    if (world) *world = pdn_comm_world(ctx);
    return PDN_OK;
}

}  // extern "C"
