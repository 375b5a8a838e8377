// cg.cuh — matrix-free conjugate-gradient solve of the pressure-Poisson system, replacing the reference's
// pre-factored Eigen::SparseLU (src/system.h:105; assembly src/waterSolver.cpp:240-261; solve :279).
//
// The reference matrix A has diag -6 in every row and +1 per in-grid neighbour (ghost pressure on all six faces
// moved to the RHS, Q10).  We solve the SPD system  M x = f,  M = -A  (M x)_c = 6 x_c - sum_{in-grid nbrs} x_n.
//
// One CG iteration = 2 kernels, 40 B/cell in fp32 (80 in fp64):
//   k_cg_dir_matvec : d' = r + beta d  (recomputed at the 6 neighbours from r,d instead of a third pass),
//                     w = M d', partial <d', w>            reads r,d  writes d',w   (16 B)
//   k_cg_update     : x += alpha d', r -= alpha w, partial <r, r>  reads x,d',r,w writes x,r (24 B)
// alpha/beta live in device memory (no host round trip); dot products accumulate in fp64 and are reduced
// warp-shuffle -> block -> fixed-order last-block pass (deterministic, no float atomics).
// Each thread owns V=4 consecutive z cells (128-bit loads/stores) when Z % 4 == 0.
#pragma once
#include "common.cuh"

namespace ink {

constexpr int kCgBlocks = 148 * 8;       // grid of the CG kernels (grid-stride)
constexpr int kCgMaxBlocks = 148 * 64;   // capacity of the per-block fp64 partials (largest grid any reducing kernel may use)

struct CgScal {
    double rr, rr_old, dw, bb, tol2;
    double rz, rz_old;   // <r, z> for preconditioned CG (z = B r); plain CG keeps rz == rr
    int iters, done;
    unsigned int ticket;
    int dist;            // multi-GPU: kernels deposit THIS RANK's partial sums in loc[]; after the NCCL all-reduce k_cg_post applies them
    double loc[4];       // [0] <d,w>  [1] <r,r>  [2] <r,z>  [3] <f,f>
    double partial[2][kCgMaxBlocks];
};

// ---- peer-memory exchanges (multi-GPU) -------------------------------------------------------------------------------------
// Device-initiated exchanges over NVLink peer memory carry the per-iteration traffic of the distributed PCG (one process per
// GPU; SURVEY 8e: y-slabs, 2 scalar all-reduces + 2 boundary-band exchanges per CG iteration).
// Why not NCCL here: each of those messages is 8 bytes .. 1 MB and there are ~5 per iteration; an NCCL call costs 15-25 us of
// launch + protocol latency regardless of size, i.e. ~110 us per iteration on top of ~120 us of compute (2 GPUs, profiles/r1_c).
// Every rank instead owns one ARENA (a single cudaMalloc, exported with cudaIpcGetMemHandle and opened by the peers once at
// start-up): peers store straight into it through NVLink and publish a sequence number; the consumer kernel spins on its LOCAL
// flag.  NCCL stays for start-up (handle all-gather) and the large once-per-step collectives (velocity all-gather, force-mask
// all-reduce), where it is at bandwidth.
// Memory-ordering contract: producer = plain stores of the payload, __threadfence_system(), then a volatile store of the
// sequence number; consumer = volatile load of the sequence number until it matches, __threadfence_system(), then the payload.
// Payload consumers in LATER kernels see it through L2 (a kernel boundary orders them after the waiting kernel).
// Boundary bands have no stand-alone put / wait kernels: the PRODUCER of a band stores it a second time into the neighbour's
// arena and its last block publishes the sequence number (P2pPush below: k_cg_init / k_cg_update for r; k_band_restore_weight in
// mg_vec.cuh for the weighted z), and the CONSUMER kernel spins on its local flag before it reads the arena (k_band_scale,
// k_band_combine).  The separate put + wait-unpack kernel pair of the first version cost two launches per exchange on the
// critical path (2 GPUs: 3.64 -> 2.99 ms per step when they were fused away).
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

// Scalar all-reduce folded into the kernel that produces the partial sums: the LAST block of a reducing kernel (cg_init,
// cg_dir_matvec, the last band_combine of an iteration) runs the mailbox exchange itself instead of a separate one-warp launch
// (two launches fewer on the critical path of every distributed PCG iteration).
struct ArFold {
    int on, rank, world, off, count, stage, precond;
    P2pPeers peers;
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

// executed by ONE FULL WARP (all 32 lanes converged)
__device__ __forceinline__ void p2p_allreduce_post_warp(CgScal* S, const P2pPeers& peers, int rank, int world, int off, int count, int stage, int precond) {
    const int lane = threadIdx.x & 31;
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
        volatile const unsigned long long* f = reinterpret_cast<const volatile unsigned long long*>(&src->seq);
        while (*f < seq) {
#if defined(__CUDA_ARCH__) && __CUDA_ARCH__ >= 700
            __nanosleep(32);
#endif
        }
        __threadfence_system();
        for (int a = 0; a < count; ++a) got[a] = *reinterpret_cast<const volatile double*>(&src->v[a]);
    }
    __syncwarp();
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
__device__ __forceinline__ void p2p_fold(CgScal* S, const ArFold& f) {   // warp 0 of the last block, after thread 0 stored S->loc
    __syncwarp();
    p2p_allreduce_post_warp(S, f.peers, f.rank, f.world, f.off, f.count, f.stage, f.precond);
}

__global__ void __launch_bounds__(32) k_p2p_allreduce_post(CgScal* S, P2pPeers peers, int rank, int world, int off, int count, int stage, int precond) {
    INK_PDL();
    if (stage != 0 && S->done) return;      // every rank holds the same `done` (same sums): all skip together
    p2p_allreduce_post_warp(S, peers, rank, world, off, count, stage, precond);
}

template <class Real, int V>
struct alignas(sizeof(Real) * V) Pack { Real a[V]; };

template <class Real, int V>
INK_HD Pack<Real, V> ldp(const Real* p) { return *reinterpret_cast<const Pack<Real, V>*>(p); }
template <class Real, int V>
INK_HD void stp(Real* p, const Pack<Real, V>& v) { *reinterpret_cast<Pack<Real, V>*>(p) = v; }

// Multi-GPU: a kernel that produces boundary rows can PUSH them straight into the neighbours' arenas (see the note above P2pMail) while it
// writes them locally, and its last block publishes the sequence number: no separate put kernel on the critical path.
// Cell offsets are relative to the first cell of this rank's slab; [0, lo_cells) goes to rank-1, [hi_start, ...) to rank+1.
struct P2pPush {
    int on;
    long long lo_cells, hi_start;
    void* dst_lo; void* dst_hi;                       // peer arena payload regions (nullptr = no neighbour on that side)
    unsigned long long* flag_lo; unsigned long long* flag_hi;
    unsigned long long seq;
};
__device__ __forceinline__ void p2p_publish(const P2pPush& p) {   // by ONE thread, after every block's stores were fenced system-wide
    __threadfence_system();
    if (p.dst_lo) *reinterpret_cast<volatile unsigned long long*>(p.flag_lo) = p.seq;
    if (p.dst_hi) *reinterpret_cast<volatile unsigned long long*>(p.flag_hi) = p.seq;
}

// block-level fp64 sum -> S->partial[slot][block]; the last block to arrive adds all partials in a fixed order
// and returns true (thread 0 of that block holds the total in `total`).
__device__ __forceinline__ bool block_sum_last(double acc, CgScal* S, int slot, double& total) {
    __shared__ double ssum[8];
    __shared__ bool is_last;
    acc = warp_sum(acc);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) ssum[wid] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double b = 0;
        for (int q = 0; q < (int)(blockDim.x >> 5); ++q) b += ssum[q];
        S->partial[slot][blockIdx.x] = b;
        __threadfence();
        const unsigned int t = atomicAdd(&S->ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return false;
    __threadfence();
    // final pass by the WHOLE last block, fixed order for a given grid size (deterministic): thread t adds partials t, t + B, ...;
    // warp shuffle; warp totals in warp order.  (One warp walking up to 9472 partials alone took microseconds.)
    double s = 0;
    for (int q = threadIdx.x; q < (int)gridDim.x; q += blockDim.x) s += __ldcg(&S->partial[slot][q]);
    s = warp_sum(s);
    __syncthreads();              // ssum is reused
    if (lane == 0) ssum[wid] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double b = 0;
        for (int q = 0; q < (int)(blockDim.x >> 5); ++q) b += ssum[q];
        total = b;
        S->ticket = 0;
    }
    return true;
}

template <class Real, int V>
__device__ __forceinline__ void vec_cell(const Grid& g, long long t, int ZV, int& i, int& j, int& k, long long& c) {
    const int kv = (int)(t % ZV);
    const long long row = t / ZV;
    i = (int)(row % g.X);
    j = g.j0 + (int)(row / g.X);
    k = kv * V;
    c = (long long)j * g.sy + (long long)i * g.sx + k;
}

// r = f - M x, <r,r>, <f,f>;  (x may be a warm start)
template <class Real, int V>
__global__ void __launch_bounds__(256) k_cg_init(Grid g, CgScal* S, const Real* __restrict__ f, const Real* __restrict__ x, Real* __restrict__ r, P2pPush push,
                                                 ArFold fold) {
    INK_PDL();
    const int ZV = g.Z / V;
    const long long nvec = (long long)(g.j1 - g.j0) * g.X * ZV;
    double arr = 0, abb = 0;
    bool pushed = false;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < nvec; t += (long long)gridDim.x * blockDim.x) {
        int i, j, k; long long c;
        vec_cell<Real, V>(g, t, ZV, i, j, k, c);
        const Pack<Real, V> xc = ldp<Real, V>(x + c), fc = ldp<Real, V>(f + c);
        Real sum[V];
#pragma unroll
        for (int a = 0; a < V; ++a) sum[a] = Real(0);
        if (i + 1 < g.X) { const Pack<Real, V> q = ldp<Real, V>(x + c + g.sx);
#pragma unroll
            for (int a = 0; a < V; ++a) sum[a] += q.a[a]; }
        if (i - 1 >= 0) { const Pack<Real, V> q = ldp<Real, V>(x + c - g.sx);
#pragma unroll
            for (int a = 0; a < V; ++a) sum[a] += q.a[a]; }
        if (j + 1 < g.Y) { const Pack<Real, V> q = ldp<Real, V>(x + c + g.sy);
#pragma unroll
            for (int a = 0; a < V; ++a) sum[a] += q.a[a]; }
        if (j - 1 >= 0) { const Pack<Real, V> q = ldp<Real, V>(x + c - g.sy);
#pragma unroll
            for (int a = 0; a < V; ++a) sum[a] += q.a[a]; }
        const Real zl = (k > 0) ? x[c - 1] : Real(0);
        const Real zr = (k + V < g.Z) ? x[c + V] : Real(0);
        Pack<Real, V> rc;
#pragma unroll
        for (int a = 0; a < V; ++a) {
            const Real left = (a > 0) ? xc.a[a > 0 ? a - 1 : 0] : zl;
            const Real right = (a < V - 1) ? xc.a[a < V - 1 ? a + 1 : 0] : zr;
            const Real mx = Real(6) * xc.a[a] - (sum[a] + left + right);
            rc.a[a] = fc.a[a] - mx;
            arr += (double)rc.a[a] * (double)rc.a[a];
            abb += (double)fc.a[a] * (double)fc.a[a];
        }
        stp<Real, V>(r + c, rc);
        if (push.on) {
            const long long off = c - (long long)g.j0 * g.sy;
            if (push.dst_lo && off < push.lo_cells) { stp<Real, V>(reinterpret_cast<Real*>(push.dst_lo) + off, rc); pushed = true; }
            if (push.dst_hi && off >= push.hi_start) { stp<Real, V>(reinterpret_cast<Real*>(push.dst_hi) + (off - push.hi_start), rc); pushed = true; }
        }
    }
    if (pushed) __threadfence_system();   // only the threads that stored to a peer pay the system-scope fence (before the block's ticket)
    double tot_rr, tot_bb = 0;
    // two reductions through the same ticket: do <f,f> first via slot 1 without the ticket (plain partials)
    {
        __shared__ double sbb[8];
        abb = warp_sum(abb);
        if ((threadIdx.x & 31) == 0) sbb[threadIdx.x >> 5] = abb;
        __syncthreads();
        if (threadIdx.x == 0) {
            double b = 0;
            for (int q = 0; q < (int)(blockDim.x >> 5); ++q) b += sbb[q];
            S->partial[1][blockIdx.x] = b;
        }
    }
    if (block_sum_last(arr, S, 0, tot_rr)) {
        if ((threadIdx.x >> 5) == 0) {
            const int lane = threadIdx.x & 31;
            double s = 0;
            for (int q = lane; q < (int)gridDim.x; q += 32) s += __ldcg(&S->partial[1][q]);
            tot_bb = warp_sum(s);
            if (lane == 0) {
                if (push.on) p2p_publish(push);
                if (S->dist) { S->loc[1] = tot_rr; S->loc[3] = tot_bb; S->loc[0] = 0; S->loc[2] = 0; }
                else {
                    S->rr = tot_rr; S->rr_old = tot_rr; S->bb = tot_bb; S->dw = 0;
                    S->rz = tot_rr; S->rz_old = tot_rr;
                    S->iters = 0;
                    S->done = (tot_rr <= S->tol2 * tot_bb) ? 1 : 0;
                }
            }
            if (fold.on) p2p_fold(S, fold);
        }
    }
}

template <class Real, int V>
__global__ void __launch_bounds__(256) k_cg_dir_matvec(Grid g, CgScal* S, const Real* __restrict__ r, const Real* __restrict__ d,
                                                       Real* __restrict__ dn, Real* __restrict__ w, int halo_lo, int halo_hi, ArFold fold) {
    INK_PDL();
    if (S->done) return;
    const Real beta = (S->iters == 0) ? Real(0) : (Real)(S->rz / S->rz_old);
    const int ZV = g.Z / V;
    const long long nvec = (long long)(g.j1 - g.j0) * g.X * ZV;
    double acc = 0;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < nvec; t += (long long)gridDim.x * blockDim.x) {
        int i, j, k; long long c;
        vec_cell<Real, V>(g, t, ZV, i, j, k, c);
        const Pack<Real, V> rc = ldp<Real, V>(r + c), dc = ldp<Real, V>(d + c);
        Real ctr[V], sum[V];
#pragma unroll
        for (int a = 0; a < V; ++a) { ctr[a] = rfma(beta, dc.a[a], rc.a[a]); sum[a] = Real(0); }
#define INK_NBR(cond, off)                                                                  \
        if (cond) {                                                                         \
            const Pack<Real, V> qr = ldp<Real, V>(r + c + (off)), qd = ldp<Real, V>(d + c + (off)); \
            _Pragma("unroll") for (int a = 0; a < V; ++a) sum[a] += rfma(beta, qd.a[a], qr.a[a]);   \
        }
        INK_NBR(i + 1 < g.X, g.sx)
        INK_NBR(i - 1 >= 0, -(long long)g.sx)
        INK_NBR(j + 1 < g.Y, g.sy)
        INK_NBR(j - 1 >= 0, -g.sy)
#undef INK_NBR
        const Real zl = (k > 0) ? rfma(beta, d[c - 1], r[c - 1]) : Real(0);
        const Real zr = (k + V < g.Z) ? rfma(beta, d[c + V], r[c + V]) : Real(0);
        Pack<Real, V> wc, dnc;
#pragma unroll
        for (int a = 0; a < V; ++a) {
            const Real left = (a > 0) ? ctr[a > 0 ? a - 1 : 0] : zl;
            const Real right = (a < V - 1) ? ctr[a < V - 1 ? a + 1 : 0] : zr;
            wc.a[a] = Real(6) * ctr[a] - (sum[a] + left + right);
            dnc.a[a] = ctr[a];
            acc += (double)ctr[a] * (double)wc.a[a];
        }
        stp<Real, V>(dn + c, dnc);
        stp<Real, V>(w + c, wc);
    }
    // multi-GPU: keep the direction current on the neighbour slabs' boundary rows too (d' = z + beta d is point-wise; z there
    // arrived by halo exchange), so the next iteration can again recompute d' at the 6 neighbours without exchanging d
    const int nh = halo_lo + halo_hi;
    if (nh) {
        const long long rowvec = (long long)g.X * ZV;
        for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < nh * rowvec; t += (long long)gridDim.x * blockDim.x) {
            const int h = (int)(t / rowvec);
            const long long rem = t - h * rowvec;
            const int j = (h == 0 && halo_lo) ? g.j0 - 1 : g.j1;
            const long long c = (long long)j * g.sy + rem * V;
            const Pack<Real, V> rc = ldp<Real, V>(r + c), dc = ldp<Real, V>(d + c);
            Pack<Real, V> o;
#pragma unroll
            for (int a = 0; a < V; ++a) o.a[a] = rfma(beta, dc.a[a], rc.a[a]);
            stp<Real, V>(dn + c, o);
        }
    }
    double tot;
    if (block_sum_last(acc, S, 0, tot)) {
        if (threadIdx.x == 0) { if (S->dist) S->loc[0] = tot; else S->dw = tot; }
        if (fold.on && threadIdx.x < 32) p2p_fold(S, fold);
    }
}

// ---- k_cg_dir_matvec through a shared-memory tile ------------------------------------------------------------------------
// The kernel above recomputes d' = r + beta d at the six neighbours of every cell: 14 vector loads per 4 cells through L1 (ncu
// r1_e: L1/TEX 58 % busy, 3 TB/s for 16 B/cell).  Here a block owns a kMvTX x kMvTY x kMvTZ tile: A) d' for the tile and its
// one-cell halo (zero outside the GRID: M only couples existing cells; rows of the neighbouring slabs are read like any other),
// once per cell, into shared memory, the tile's own d' also to global memory; B) w = 6 d' - (six neighbours) from shared
// memory, <d', w>.  1.6 x 8 B/cell from L2 instead of 7 x 8 B/cell through L1.  Z % 4 == 0.
constexpr int kMvTX = 8, kMvTY = 8, kMvTZ = 64;
constexpr int kMvRows = (kMvTX + 2) * (kMvTY + 2);
constexpr int kMvRow = kMvTZ + 8;                        // [3 pad][k = -1][TZ cells][k = TZ][3 pad]
template <class Real>
constexpr size_t cg_matvec_tile_smem() { return (size_t)kMvRows * kMvRow * sizeof(Real); }

template <class Real>
__global__ void __launch_bounds__(256) k_cg_dir_matvec_tile(Grid g, CgScal* S, const Real* __restrict__ r, const Real* __restrict__ d, Real* __restrict__ dn,
                                                            Real* __restrict__ w, int halo_lo, int halo_hi, ArFold fold, int ntx, int nty, int ntz) {
    INK_PDL();
    if (S->done) return;
    extern __shared__ __align__(16) unsigned char mv_smem_raw[];
    Real* xs = reinterpret_cast<Real*>(mv_smem_raw);
    const Real beta = (S->iters == 0) ? Real(0) : (Real)(S->rz / S->rz_old);
    const int tid = threadIdx.x;
    const int sxf = g.sx, syf = (int)g.sy;                   // launcher guarantees X*Y*Z < 2^31
    double acc = 0;
    for (int tile = blockIdx.x; tile < ntx * nty * ntz; tile += gridDim.x) {
        const int tz_ = tile % ntz, txy = tile / ntz, tx_ = txy % ntx, ty_ = txy / ntx;
        const int fx0 = tx_ * kMvTX, fy0 = g.j0 + ty_ * kMvTY, fz0 = tz_ * kMvTZ;
        // ---- A: d' on the tile + halo rows (corner rows are nobody's face neighbours)
        for (int q = tid; q < kMvRows * (kMvTZ / 4); q += 256) {
            const int zq = q % (kMvTZ / 4), row = q / (kMvTZ / 4), lx = row % (kMvTX + 2), ly = row / (kMvTX + 2);
            if ((lx == 0 || lx == kMvTX + 1) && (ly == 0 || ly == kMvTY + 1)) continue;
            const int fi = fx0 - 1 + lx, fj = fy0 - 1 + ly, fk = fz0 + 4 * zq;
            Pack<Real, 4> v;
            v.a[0] = v.a[1] = v.a[2] = v.a[3] = Real(0);
            if (fi >= 0 && fi < g.X && fj >= 0 && fj < g.Y && fk < g.Z) {
                const int c = fj * syf + fi * sxf + fk;
                const Pack<Real, 4> rc = ldp<Real, 4>(r + c), dc = ldp<Real, 4>(d + c);
#pragma unroll
                for (int a = 0; a < 4; ++a) v.a[a] = rfma(beta, dc.a[a], rc.a[a]);
                if (lx >= 1 && lx <= kMvTX && ly >= 1 && ly <= kMvTY && fj < g.j1) stp<Real, 4>(dn + c, v);   // this rank's own cells
            }
            stp<Real, 4>(xs + row * kMvRow + 4 + 4 * zq, v);
        }
        for (int q = tid; q < kMvTX * kMvTY * 2; q += 256) {   // the two z-halo cells of the interior rows
            const int side = q & 1, col = q >> 1, lx = 1 + col % kMvTX, ly = 1 + col / kMvTX;
            const int fi = fx0 - 1 + lx, fj = fy0 - 1 + ly, fk = side ? fz0 + kMvTZ : fz0 - 1;
            Real v = Real(0);
            if (fi < g.X && fj < g.Y && fk >= 0 && fk < g.Z) { const int c = fj * syf + fi * sxf + fk; v = rfma(beta, d[c], r[c]); }
            xs[(ly * (kMvTX + 2) + lx) * kMvRow + (side ? 4 + kMvTZ : 3)] = v;
        }
        __syncthreads();
        // ---- B: w = M d' on the tile
        for (int q = tid; q < kMvTX * kMvTY * (kMvTZ / 4); q += 256) {
            const int zq = q % (kMvTZ / 4), col = q / (kMvTZ / 4), lx = 1 + col % kMvTX, ly = 1 + col / kMvTX;
            const int fi = fx0 - 1 + lx, fj = fy0 - 1 + ly, fk = fz0 + 4 * zq;
            if (fi >= g.X || fj >= g.j1 || fk >= g.Z) continue;
            const Real* c = xs + (ly * (kMvTX + 2) + lx) * kMvRow + 4 + 4 * zq;
            const Pack<Real, 4> xc = ldp<Real, 4>(c), xl = ldp<Real, 4>(c - kMvRow), xr = ldp<Real, 4>(c + kMvRow);
            const Pack<Real, 4> yl = ldp<Real, 4>(c - (kMvTX + 2) * kMvRow), yr = ldp<Real, 4>(c + (kMvTX + 2) * kMvRow);
            const Real zl = c[-1], zr = c[4];
            Pack<Real, 4> wc;
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const Real lo = (a == 0) ? zl : xc.a[a > 0 ? a - 1 : 0], hi = (a == 3) ? zr : xc.a[a < 3 ? a + 1 : 3];
                // same association as k_cg_dir_matvec: ((x+ + x-) + (y+ + y-) accumulated in that order) + left + right
                const Real sum = ((xr.a[a] + xl.a[a]) + yr.a[a]) + yl.a[a];
                wc.a[a] = Real(6) * xc.a[a] - (sum + lo + hi);
                acc += (double)xc.a[a] * (double)wc.a[a];
            }
            stp<Real, 4>(w + (fj * syf + fi * sxf + fk), wc);
        }
        __syncthreads();
    }
    // multi-GPU: keep the direction current on the neighbour slabs' boundary rows too (see k_cg_dir_matvec)
    const int nh = halo_lo + halo_hi;
    if (nh) {
        const int ZV = g.Z / 4;
        const long long rowvec = (long long)g.X * ZV;
        for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < nh * rowvec; t += (long long)gridDim.x * blockDim.x) {
            const int h = (int)(t / rowvec);
            const long long rem = t - h * rowvec;
            const int j = (h == 0 && halo_lo) ? g.j0 - 1 : g.j1;
            const long long c = (long long)j * g.sy + rem * 4;
            const Pack<Real, 4> rc = ldp<Real, 4>(r + c), dc = ldp<Real, 4>(d + c);
            Pack<Real, 4> o;
#pragma unroll
            for (int a = 0; a < 4; ++a) o.a[a] = rfma(beta, dc.a[a], rc.a[a]);
            stp<Real, 4>(dn + c, o);
        }
    }
    double tot;
    if (block_sum_last(acc, S, 0, tot)) {
        if (threadIdx.x == 0) { if (S->dist) S->loc[0] = tot; else S->dw = tot; }
        if (fold.on && threadIdx.x < 32) p2p_fold(S, fold);
    }
}

// PRECOND: z = B r is produced by the V-cycle that follows (mg.cuh), which also sets rz / rz_old
template <class Real, int V, bool PRECOND>
__global__ void __launch_bounds__(256) k_cg_update(Grid g, CgScal* S, Real* __restrict__ x, Real* __restrict__ r, const Real* __restrict__ d,
                                                   const Real* __restrict__ w, P2pPush push) {
    INK_PDL();
    if (S->done) return;
    const Real alpha = (Real)(S->rz / S->dw);
    const int ZV = g.Z / V;
    const long long nvec = (long long)(g.j1 - g.j0) * g.X * ZV;
    const long long base = (long long)g.j0 * g.sy;
    double acc = 0;
    bool pushed = false;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < nvec; t += (long long)gridDim.x * blockDim.x) {
        const long long c = base + t * V;
        Pack<Real, V> xc = ldp<Real, V>(x + c), rc = ldp<Real, V>(r + c);
        const Pack<Real, V> dc = ldp<Real, V>(d + c), wc = ldp<Real, V>(w + c);
#pragma unroll
        for (int a = 0; a < V; ++a) {
            xc.a[a] = rfma(alpha, dc.a[a], xc.a[a]);
            rc.a[a] = rfma(-alpha, wc.a[a], rc.a[a]);
            acc += (double)rc.a[a] * (double)rc.a[a];
        }
        stp<Real, V>(x + c, xc);
        stp<Real, V>(r + c, rc);
        if (push.on) {
            const long long off = t * V;
            if (push.dst_lo && off < push.lo_cells) { stp<Real, V>(reinterpret_cast<Real*>(push.dst_lo) + off, rc); pushed = true; }
            if (push.dst_hi && off >= push.hi_start) { stp<Real, V>(reinterpret_cast<Real*>(push.dst_hi) + (off - push.hi_start), rc); pushed = true; }
        }
    }
    if (pushed) __threadfence_system();   // only the threads that stored to a peer pay the system-scope fence (before the block's ticket)
    double tot;
    if (block_sum_last(acc, S, 1, tot)) {
        if (threadIdx.x == 0) {
            if (push.on) p2p_publish(push);
            if (S->dist) S->loc[1] = tot;
            else {
                S->rr_old = S->rr;
                S->rr = tot;
                if (!PRECOND) { S->rz_old = S->rz; S->rz = tot; }
                S->iters += 1;
                if (tot <= S->tol2 * S->bb) S->done = 1;
            }
        }
    }
}

// Small device -> pinned-host readback WITHOUT the copy engine: one warp stores the words straight into mapped host memory
// (UVA: a cudaMallocHost pointer is valid on the device).  The step's host-visible scalars (dt, CG convergence head, out-of-
// grid particle count) used to go through cudaMemcpyAsync on the compute stream, i.e. through the D2H copy engine, where they
// queued behind a 120 MB particle download running on the output copy stream: the whole step then waited for that copy.
__global__ void k_copy_words(uint32_t* __restrict__ dst_host, const uint32_t* __restrict__ src, int nwords) {
    INK_PDL();
    for (int t = threadIdx.x; t < nwords; t += blockDim.x) dst_host[t] = src[t];
    __threadfence_system();
}

// Multi-GPU only: after ncclAllReduce(sum) of S->loc, one thread applies the global sums exactly like the fused single-GPU
// epilogues do.  stage 0: after k_cg_init; 1: after k_cg_dir_matvec; 2: end of iteration (rr from k_cg_update, rz from the
// V-cycle when preconditioned, else rz = rr); 3: <r,z> of the very first V-cycle.
__global__ void k_cg_post(CgScal* S, int stage, int precond) {
    INK_PDL();
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (stage == 0) {
        S->rr = S->loc[1]; S->rr_old = S->loc[1]; S->bb = S->loc[3]; S->dw = 0;
        S->rz = S->loc[1]; S->rz_old = S->loc[1];
        S->iters = 0;
        S->done = (S->loc[1] <= S->tol2 * S->loc[3]) ? 1 : 0;
    } else if (S->done) {
        return;
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

}  // namespace ink
