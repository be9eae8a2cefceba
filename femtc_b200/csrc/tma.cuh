// PTX wrappers shared by the tiled (slab) and column kernels: mbarriers and 1-D TMA bulk copies
// (cp.async.bulk; SASS: UBLKCP / SYNCS).  sm_100a only.
#pragma once

#include <cstdint>

namespace femtc
{
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Generic-proxy reads of a staging buffer followed by a bulk copy (async proxy) that refills it: a write-after-read
// hazard that __syncwarp() and a CTA fence do not close (measured in round 1: slab mass kernel, fp32 n1 = 8, 48 of 399
// launches wrong without a fence, ~1 in 500 with one; tools/probe/repeat_probe.py) -- a refill served from L2 can
// overtake shared-memory loads still queued behind bank conflicts.  Two rules close it, neither depends on where the
// compiler places instructions:
//  (1) buffers whose values go straight into a shared-memory store or a block barrier (U, detJ staging) are refilled
//      only after that store / barrier: the store cannot issue before the loads have returned;
//  (2) ring slots whose values are consumed into registers only (the geometry rings) are handed back through a
//      per-slot `empty` mbarrier, the consumer-release pattern of every TMA pipeline: each reading lane executes
//      mbarrier.arrive (release.cta: the lane's prior shared-memory reads are performed before the arrival, and the
//      arrival is queued behind them in the same in-order shared-memory pipe), and the lanes that issue the refill
//      wait (acquire) for the phase, i.e. for all arrivals, before cp.async.bulk.  slot_release_and_acquire() below.
//  (2') block-wide rings filled by a producer warp (column_kernels.cu, PW; the n1 >= 7 stiffness defaults) hand a slot back
//      with `__syncwarp(); if (lane == 0) mbar_arrive(empty)` and an `empty` count of one arrival per compute warp.  That
//      is rule 2 with one arrival standing for the warp, and it is sufficient there because the refill is issued by
//      ANOTHER warp behind an acquire: (a) __syncwarp orders every lane's earlier shared-memory reads before anything
//      lane 0 does after it, (b) lane 0's arrive is a release and release is cumulative, so those reads are performed
//      before the producer's try_wait (acquire) can observe the completed phase, and only then does the producer issue
//      the bulk copy; (c) in the pipe, the warp's loads are warp-wide instructions issued before the arrive instruction
//      of the same converged warp, so the arrival cannot be queued ahead of another lane's load.  The round-1 failure had
//      no such hand-over (loads, __syncwarp, fence and the refill all in one warp, nothing waiting on an arrival).
//      The form is the consumer release of CUTLASS's TMA pipelines (one elected arrive per consumer warp).  Evidence for
//      these kernels specifically: tests/test_gpu_parity.py::test_repeated_launches_on_cache_resident_fields_are_bit_identical runs the
//      default fp32 REF_EXACT stiffness kernels at n1 = 7 and 9 (block-wide rings) 1000 times against their first result.
#ifndef FEMTC_RELEASE_LANES
#define FEMTC_RELEASE_LANES 32 // arrivals per warp on an `empty` barrier (32: every reading lane; 1: lane 0 after __syncwarp, A/B only)
#endif
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
  asm volatile("{\n"
               ".reg .pred P1;\n"
               "LAB_WAIT:\n"
               "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
               "@P1 bra DONE;\n"
               "bra LAB_WAIT;\n"
               "DONE:\n"
               "}" ::"r"(bar),
               "r"(parity)
               : "memory");
}
// Every lane of a (converged) warp: "my reads of this ring slot are done", then wait until all lanes have said so.
// On return the calling lane may refill the slot with a bulk copy.  `parity` = (number of earlier uses of the slot) & 1.
__device__ __forceinline__ void slot_release_and_acquire(uint32_t bar_empty, uint32_t parity, int lane)
{
#if FEMTC_RELEASE_LANES == 32
  (void)lane;
  mbar_arrive(bar_empty);
#else
  __syncwarp();
  if (lane == 0)
    mbar_arrive(bar_empty);
#endif
  mbar_wait(bar_empty, parity);
}
// global -> shared bulk copy (TMA), completion signalled on an mbarrier
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// shared -> global bulk copy (TMA), tracked by the issuing thread's bulk async-group
__device__ __forceinline__ void bulk_s2g(void* dst, uint32_t src, uint32_t bytes)
{
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
} // namespace femtc
