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

// (P2pMail / P2pHead / P2pPeers, the scalar all-reduce and its fold-in form live in cg.cuh, next to the kernels that end with them)

// ---- boundary bands ------------------------------------------------------------------------------------------------------
// There are no stand-alone put / wait kernels: the PRODUCER of a band stores it a second time into the neighbour's arena
// and its last block publishes the sequence number (P2pPush in cg.cuh: k_cg_init / k_cg_update for r; k_band_restore_weight
// in mg_vec.cuh for the weighted z), and the CONSUMER kernel spins on its local flag before it reads the arena
// (k_band_scale, k_band_combine).  The separate k_p2p_put + k_p2p_wait_unpack pair of the first version cost two launches
// per exchange on the critical path (2 GPUs: 3.64 -> 2.99 ms per step when they were fused away).

}  // namespace ink
