// TMA (cp.async.bulk.tensor) + mbarrier helpers shared by the kernels that stage image windows through the tensor memory accelerator,
// and the per-level tensor-map set they take as a __grid_constant__ parameter.
//
// Every level plane of a pass workspace (pyramid or blurred) is described as a 3-D u8 tensor [frame][plane row][byte column] that starts at
// the plane ALLOCATION (apron rows and the 16-byte row lead included), so window coordinates are never negative:
//   tensor column = level column + 16, tensor row = level row + ORB_APRON_ROWS.
// The innermost coordinate of a box (its first byte column) must be a multiple of 16: an unaligned start raises an illegal-instruction
// fault on this driver (tests/cuda/tma_probe.cu). Bytes of a box that fall outside the tensor arrive as zeros. The box size is part of the descriptor (orb_build_level_maps).
#pragma once
#include <cuda.h>

#include "orb_common.cuh"

struct OrbLevelMaps { CUtensorMap m[ORB_MAX_LEVELS]; };   // 128 B each, 64-byte aligned
#define ORB_LEVEL_MAPS_BYTES (128 * ORB_MAX_LEVELS)

#ifdef __CUDACC__
__device__ __forceinline__ unsigned orb_smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void orb_mbar_init(unsigned bar, unsigned arrivals) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(arrivals));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void orb_mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// one try_wait: suspends the thread for a hardware time slice at most; returns 1 when the phase with this parity has completed
__device__ __forceinline__ unsigned orb_mbar_try_wait(unsigned bar, unsigned parity) {
  unsigned done;
  asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
               : "=r"(done)
               : "r"(bar), "r"(parity)
               : "memory");
  return done;
}
// bounded wait: a lost copy becomes a reported error (returns 0) instead of a hang
__device__ __forceinline__ unsigned orb_mbar_wait(unsigned bar, unsigned parity) {
  unsigned done;
  int spins = 0;
  do { done = orb_mbar_try_wait(bar, parity); } while (!done && ++spins < (1 << 20));
  return done;
}
// box load global -> shared (dst 128-byte aligned), completion counted in bytes on the mbarrier
__device__ __forceinline__ void orb_tma_load_3d(unsigned dst, const CUtensorMap* map, int c0, int c1, int c2, unsigned bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
               "l"(reinterpret_cast<unsigned long long>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
               : "memory");
}
// plain bulk copy global -> shared (both 16-byte aligned, bytes % 16 == 0), completion counted on the same kind of mbarrier
__device__ __forceinline__ void orb_bulk_load_1d(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
#endif
