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
    unsigned int ticket[4];                 // LOCAL: block tickets of the put kernels
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

// ---- boundary-band put: copy up to two contiguous ranges into the two neighbours' arenas, then publish `seq` ------------
struct P2pPut {
    const void* src[2];       // local source (nullptr = no neighbour on that side)
    void* dst[2];             // peer arena address of the payload region
    unsigned long long* flag[2];   // peer arena address of the flag to publish
    long long n16[2];         // payload size in 16-byte units
};

__global__ void __launch_bounds__(256) k_p2p_put(const CgScal* S, P2pPut d, unsigned long long seq, unsigned int* ticket) {
    INK_PDL();
    if (S->done) return;
    for (int b = 0; b < 2; ++b) {
        if (!d.src[b]) continue;
        const uint4* __restrict__ s = reinterpret_cast<const uint4*>(d.src[b]);
        uint4* __restrict__ t = reinterpret_cast<uint4*>(d.dst[b]);
        for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < d.n16[b]; q += (long long)gridDim.x * blockDim.x) t[q] = s[q];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int tk = atomicAdd(ticket, 1u);
        if (tk == gridDim.x - 1) {          // every block's stores are fenced: publish
            *ticket = 0;
            __threadfence_system();
            for (int b = 0; b < 2; ++b)
                if (d.src[b]) *reinterpret_cast<volatile unsigned long long*>(d.flag[b]) = seq;
        }
    }
}

// ---- wait for both neighbours' bands of sequence `seq`, optionally unpack them into their rows ------------------------
struct P2pGet {
    const unsigned long long* flag[2];   // LOCAL flags (nullptr = no neighbour on that side)
    const void* src[2];                  // local arena payload
    void* dst[2];                        // where to unpack (nullptr = leave in the arena)
    long long n16[2];
};

__global__ void __launch_bounds__(256) k_p2p_wait_unpack(const CgScal* S, P2pGet d, unsigned long long seq) {
    INK_PDL();
    if (S->done) return;
    if (threadIdx.x == 0)
        for (int b = 0; b < 2; ++b)
            if (d.flag[b]) p2p_spin(reinterpret_cast<const volatile unsigned long long*>(d.flag[b]), seq);
    __syncthreads();
    for (int b = 0; b < 2; ++b) {
        if (!d.flag[b] || !d.dst[b]) continue;
        const uint4* s = reinterpret_cast<const uint4*>(d.src[b]);
        uint4* t = reinterpret_cast<uint4*>(d.dst[b]);
        for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < d.n16[b]; q += (long long)gridDim.x * blockDim.x) {
            uint4 v;
            asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(s + q));   // L2: the peer wrote it
            t[q] = v;
        }
    }
}

}  // namespace ink
