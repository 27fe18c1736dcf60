// pdn_peer.cuh -- multi-GPU exchanges as peer-memory stores over NVLink / NVSwitch.
//
// NCCL's small-message latency (measured: ~30 us per collective inside the CUDA graph, three per
// iteration) is larger than a whole SpMV on a 1M-row partition.  The three exchanges of an iteration
// carry a few KB, so here every rank maps one "arena" of every peer (CUDA IPC) and
//   * halo:       the owner of an interface entry stores it straight into the peers' copies of the
//                 vector (same global index: no pack, no unpack), then raises a flag on each peer;
//   * dot sums:   each rank stores its 3 partial sums into a slot on every rank, every rank adds
//                 the slots in rank order -- identical bits everywhere, no reduction tree;
//   * coarse r_c: the same with nc 64-bit fixed-point partial sums (exact, order-independent).
// Flags carry an epoch (one counter per exchange type, incremented identically on all ranks), payload
// slots are double-buffered by epoch parity.  A rank can run at most one epoch ahead of its slowest
// peer because each epoch consumes every peer's contribution to the previous one.
// Every spin has a clock64() deadline (~2 s): a lost peer becomes PDN_ECUDA, never a hung GPU.
//
// Included by pdn_kernels.cuh (namespace pdnk).
#pragma once

// ---- halo ------------------------------------------------------------------------------------------
// vec_off: word offset of the vector inside every arena.  send_idx holds, peer after peer, the global
// indices this rank owns and peer k reads (send_off / send_cnt per peer).
__global__ void __launch_bounds__(256) halo_push_kernel(PeerLink L, HaloSend h, size_t vec_off,
                                                        const int32_t *__restrict__ send_idx, unsigned int *ticket) {
    const unsigned long long *mine = L.arena[L.me] + vec_off;
    for (int k = 0; k < L.world; k++) {
        if (k == L.me || h.cnt[k] == 0) continue;
        unsigned long long *dst = L.arena[k] + vec_off;
        const int32_t *idx = send_idx + h.off[k];
        for (int32_t i = blockIdx.x * 256 + threadIdx.x; i < h.cnt[k]; i += gridDim.x * 256) {
            const int32_t g = idx[i];
            dst[g] = mine[g];
        }
    }
    __threadfence_system();
    __shared__ bool last;
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    if (threadIdx.x == 0) *ticket = 0u;
    __threadfence_system();
    const unsigned long long e = L.epoch[0] + 1;
    if (threadIdx.x < L.world && threadIdx.x != L.me && h.cnt[threadIdx.x] > 0)
        st_release_sys(L.arena[threadIdx.x] + PA_HALO_FLAG + L.me, e);
    __syncthreads();
    if (threadIdx.x == 0) L.epoch[0] = e;
}
// one warp: lane k waits for peer k's entries of the current epoch
__global__ void halo_wait_kernel(PeerLink L, HaloSend h) {
    const int k = threadIdx.x;
    if (k < L.world && k != L.me && h.recv_cnt[k] > 0)
        peer_wait(L.arena[L.me] + PA_HALO_FLAG + k, L.epoch[0], L.status, L.epoch + 4);
}

// ---- dot products ----------------------------------------------------------------------------------
// sc->part[0..2] of every rank -> summed in rank order on every rank, then the scalar recurrence.
// WHICH as in scalar_kernel.
template <int WHICH>
__global__ void peer_reduce_kernel(PeerLink L, Scal *sc, double *out) {
    double acc[3] = {sc->part[0], sc->part[1], sc->part[2]};      // one warp; lane 0's copy is the one that counts
    const bool ok = peer_allreduce3(L, acc);
    if (threadIdx.x != 0) return;
    if (!ok) { sc->done = 3; sc->breakdown = 1; return; }
    sc->part[0] = acc[0]; sc->part[1] = acc[1]; sc->part[2] = acc[2];
    if (WHICH != RESID && WHICH != 7 && WHICH != 6 && sc->done) return;
    if (WHICH == 5) bicg_x_finalize(acc, sc);
    else if (WHICH == 6) bicg_init_finalize(acc, sc);
    else if (WHICH == 7) *out = acc[0];
    else spmv_finalize<WHICH>(acc, sc);
}

// ---- coarse residual -----------------------------------------------------------------------------------
// the active slot of rc_fix (this rank's fixed-point partial sums, nc values; slot < 0: iteration parity)
// -> exact sum over the ranks, in place
__global__ void __launch_bounds__(256) peer_sum_i64_kernel(PeerLink L, const Scal *__restrict__ sc, long long *rc_fix, int32_t nc,
                                                           int rc_slot, const int32_t *arange) {
    if (sc->done) return;    // identical on every rank
    long long *rc = rc_fix + (size_t)(rc_slot < 0 ? (sc->iter & 1) : rc_slot) * nc;
    peer_sum_i64_cta(L, rc, nc, arange);
}
