// pipe.cuh -- sm_100a plumbing shared by the CTA-tile kernels: mbarriers, 1-D
// TMA bulk copies (cp.async.bulk, SASS UBLKCP) and named barriers.
//
// Kernel anatomy used by every codec kernel built on this file:
//   * persistent CTAs (cooperative launch, all resident), static round-robin
//     over tiles: CTA b handles tiles b, b + grid, b + 2 grid, ...  so tile g
//     only ever depends on tiles < g, which are running or done;
//   * warp roles: consumer warps (the codec math), one producer warp (fetches
//     the tile's metadata and issues the bulk copies that stage the tile's
//     input in a shared-memory ring, several tiles ahead of the consumers), one
//     scan warp (decoupled look-back over the tile aggregates, off the
//     consumers' critical path);
//   * the consumers run the COUNT phase of tile i + 1 (aggregate only) before
//     the EMIT phase of tile i, so the prefix of a tile has a whole tile time
//     to resolve and nobody spins on a predecessor that is still loading.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace pg {

__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
	return (uint32_t) __cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
// make the inits visible to the async proxy (TMA complete_tx) before first use
__device__ __forceinline__ void mbar_fence_init()
{
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
		     : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity)
{
	uint32_t ok;
	asm volatile(
		"{\n\t.reg .pred p;\n\t"
		"mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
		"selp.u32 %0, 1, 0, p;\n\t}"
		: "=r"(ok)
		: "r"(smem_u32(bar)), "r"(parity), "r"(4000u)	// suspend-time hint (ns): idle roles stay off the issue ports
		: "memory");
	return ok != 0;
}
// Block until the phase with the given parity has completed.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
	while (!mbar_try_wait(bar, parity))
		__nanosleep(200);	// waiting roles must not take issue slots from working ones
}

// 1-D bulk copy global -> shared, completion signalled on `bar` (complete_tx).
// dst, src and bytes must be multiples of 16.
__device__ __forceinline__ void tma_load_1d(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
	asm volatile(
		"cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
		::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar))
		: "memory");
}

// barrier among `nthreads` threads of the CTA (a multiple of 32), id 1..15
__device__ __forceinline__ void named_bar_sync(int id, int nthreads)
{
	asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

}  // namespace pg
