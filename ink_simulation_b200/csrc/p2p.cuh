// p2p.cuh — device-initiated exchanges over NVLink peer memory for the per-iteration traffic of the distributed PCG
// (one process per GPU; SURVEY 8e: y-slabs, 2 scalar all-reduces + 2 boundary-band exchanges per CG iteration).
//
// Why not NCCL here: each of those messages is 8 bytes .. 1 MB and there are ~5 per iteration; an NCCL call costs
// 15-25 us of launch + protocol latency regardless of size, i.e. ~110 us per iteration on top of ~120 us of compute
// (2 GPUs, profiles/r1_c).  Every rank instead owns one ARENA (a single cudaMalloc, exported with cudaIpcGetMemHandle
// and opened by the peers once at start-up): peers store straight into it through NVLink and publish a sequence number;
// the consumer kernel spins on its LOCAL flag.  NCCL stays for start-up (handle all-gather) and the large once-per-step
// collectives (velocity all-gather, force-mask all-reduce), where it is at bandwidth.
//
// Memory-ordering contract: producer = plain stores of the payload, __threadfence_system(), then a volatile store of the
// sequence number; consumer = volatile load of the sequence number until it matches, __threadfence_system(), then the
// payload.  Payload consumers in LATER kernels see it through L2 (a kernel boundary orders them after the waiting kernel).
#pragma once
#include "cg.cuh"
#include "common.cuh"

namespace ink {

constexpr int kP2pMaxRanks = 16;

struct alignas(64) P2pMail {
    double v[4];
    unsigned long long seq;
    unsigned long long pad[3];
};

// the head of every rank's arena (payload regions follow at byte offsets the host computes, identical on all ranks)
struct P2pHead {
    P2pMail mail[2][kP2pMaxRanks];          // [slot][source rank]
    unsigned long long flag[4][2];          // [channel][side 0 = from rank-1, 1 = from rank+1] last sequence delivered
    unsigned long long mail_seq;            // LOCAL: all-reduces this rank has taken part in (kept on the device so that
                                            // calls skipped after convergence do not advance it)
    unsigned int ticket[4];                 // LOCAL: block tickets of the pushing band kernels ([2 + side])
};

struct P2pPeers {
    unsigned char* base[kP2pMaxRanks];      // arena of every rank as mapped in THIS process (own arena at [rank])
};

// ---- scalar all-reduce (sum of `count` <= 4 doubles starting at S->loc[off]) fused with the CG scalar epilogue ---------
// One warp: lane p < world stores this rank's partials into rank p's mailbox, then waits for rank p's partials in the
// local mailbox; lane 0 adds them in rank order (identical result on every rank) and applies k_cg_post's logic.
__device__ __forceinline__ void cg_post_apply(CgScal* S, int stage, int precond) {
    if (stage == 0) {
        S->rr = S->loc[1]; S->rr_old = S->loc[1]; S->bb = S->loc[3]; S->dw = 0;
        S->rz = S->loc[1]; S->rz_old = S->loc[1];
        S->iters = 0;
        S->done = (S->loc[1] <= S->tol2 * S->loc[3]) ? 1 : 0;
    } else if (stage == 1) {
        S->dw = S->loc[0];
    } else if (stage == 2) {
        S->rr_old = S->rr; S->rr = S->loc[1];
        S->rz_old = S->rz; S->rz = precond ? S->loc[2] : S->loc[1];
        S->iters += 1;
        if (S->loc[1] <= S->tol2 * S->bb) S->done = 1;
    } else {
        S->rz_old = S->loc[2]; S->rz = S->loc[2];
    }
}

__global__ void __launch_bounds__(32) k_p2p_allreduce_post(CgScal* S, P2pPeers peers, int rank, int world, int off, int count, int stage, int precond) {
    INK_PDL();
    if (stage != 0 && S->done) return;      // every rank holds the same `done` (same sums): all skip together
    const int lane = threadIdx.x;
    P2pHead* me = reinterpret_cast<P2pHead*>(peers.base[rank]);
    const unsigned long long seq = me->mail_seq + 1;
    const int slot = (int)(seq & 1ull);
    if (lane < world) {
        P2pMail* dst = &reinterpret_cast<P2pHead*>(peers.base[lane])->mail[slot][rank];
        for (int a = 0; a < count; ++a) dst->v[a] = S->loc[off + a];
        __threadfence_system();
        *reinterpret_cast<volatile unsigned long long*>(&dst->seq) = seq;
    }
    double got[4] = {0, 0, 0, 0};
    if (lane < world) {
        const P2pMail* src = &me->mail[slot][lane];
        p2p_spin(reinterpret_cast<const volatile unsigned long long*>(&src->seq), seq);
        for (int a = 0; a < count; ++a) got[a] = *reinterpret_cast<const volatile double*>(&src->v[a]);
    }
    for (int a = 0; a < count; ++a) {       // fixed rank order 0, 1, .. world-1 on every rank
        double tot = 0;
        for (int p = 0; p < world; ++p) tot += __shfl_sync(0xffffffffu, got[a], p);
        if (lane == 0) S->loc[off + a] = tot;
    }
    __syncwarp();
    if (lane == 0) {
        me->mail_seq = seq;
        cg_post_apply(S, stage, precond);
    }
}

// ---- boundary bands ------------------------------------------------------------------------------------------------------
// There are no stand-alone put / wait kernels: the PRODUCER of a band stores it a second time into the neighbour's arena
// and its last block publishes the sequence number (P2pPush in cg.cuh: k_cg_init / k_cg_update for r; k_band_restore_weight
// in mg_vec.cuh for the weighted z), and the CONSUMER kernel spins on its local flag before it reads the arena
// (k_band_scale, k_band_combine).  The separate k_p2p_put + k_p2p_wait_unpack pair of the first version cost two launches
// per exchange on the critical path (2 GPUs: 3.64 -> 2.99 ms per step when they were fused away).

}  // namespace ink
