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
// Generic-proxy reads of a staging buffer followed by a bulk copy that refills it: when the values read are not yet
// consumed (no store or barrier depends on them), nothing orders the reads before the async-proxy write, and a refill
// served from L2 can overtake shared-memory loads still queued behind bank conflicts (measured: slab mass kernel,
// fp32 n1 = 8, 48 of 399 launches wrong; tools/probe/repeat_probe.py).  Every reading lane calls this before the
// __syncwarp() / barrier that precedes the refill; the CTA-scope fence waits for the lane's outstanding loads.
// Not needed where the values are consumed first (a store or block barrier depends on them) or where the slot is
// handed to a producer warp by mbarrier.arrive (release; the canonical consumer-release pattern).
#ifndef FEMTC_REFILL_FENCE
#define FEMTC_REFILL_FENCE 1
#endif
__device__ __forceinline__ void staging_reads_done()
{
#if FEMTC_REFILL_FENCE
  __threadfence_block();
#endif
}
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
