// tcgen05 / TMEM plumbing for sm_100a: PTX wrappers, shared-memory matrix descriptors and the
// "operand image" layout the split-precision GEMM (gemm3x_kernel.cuh) reads.
//
// Operand image.  A K-major matrix Mtx[R][K] (f32, R a multiple of 8, K a multiple of 32) is
// stored as  [K/32 k-blocks][R/8 row groups][8 rows][128 bytes]  where each 1024-byte
// (8 row x 32 float) atom is already in the 128-byte-swizzled order the tensor core expects
// (16-byte chunk index XOR row-in-atom).  For a fixed k-block any run of consecutive row
// groups is contiguous, so ONE 1-D bulk copy (cp.async.bulk) brings a (rows x 32) operand
// tile into shared memory ready for tcgen05.mma -- no tensor map, no driver API.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace b200mc {
namespace umma {

constexpr int kBlockK = 32;      // f32 elements per k-block = one 128-byte swizzle row
constexpr int kAtomBytes = 1024; // 8 rows x 128 B

// Byte offset of element (r, k) inside an operand image with R rows.
__host__ __device__ inline long image_offset(int R, int r, int k) {
  const int kb = k >> 5, kk = k & 31;
  const int ra = r >> 3, rr = r & 7;
  return ((long)kb * (R >> 3) + ra) * kAtomBytes + rr * 128 + (((kk >> 2) ^ rr) << 4) + ((kk & 3) << 2);
}

// x = hi + lo with hi, lo representable in TF32 (10 explicit mantissa bits, round to nearest):
// hi*hi' + hi*lo' + lo*hi' reproduces the f32 product to ~2^-22 ("3xTF32").
__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
  uint32_t h;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
  hi = __uint_as_float(h);
  const float rest = x - hi;  // exact
  uint32_t l;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(rest));
  lo = __uint_as_float(l);
}

__device__ __forceinline__ uint32_t smem_addr(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// Shared-memory matrix descriptor, K-major, SWIZZLE_128B, tile = [rows][128 B] with 8-row atoms
// 1024 B apart (CUTLASS cute/arch/mma_sm100_desc.hpp SmemDescriptor: start >> 4 in [0,14),
// LBO >> 4 in [16,30) (unused for swizzled K-major, 1), SBO >> 4 in [32,46), version 1 in
// [46,48), layout type 2 = SWIZZLE_128B in [61,64)).  The tile must be 1024-byte aligned.
__device__ __forceinline__ uint64_t make_desc_k_sw128(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | (1ull << 16) | ((uint64_t)(kAtomBytes >> 4) << 32) |
         (1ull << 46) | (2ull << 61);
}

// Instruction descriptor for kind::tf32, f32 accumulate, both operands K-major
// (InstrDescriptor: c_format [4,6) = 1 (F32), a_format [7,10) = b_format [10,13) = 2 (TF32),
// n_dim [17,23) = N >> 3, m_dim [24,29) = M >> 4).
__host__ __device__ constexpr uint32_t make_idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {  // one full warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(dst_smem)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {  // one full warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void fence_after_sync() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// D[tmem] (+)= A[smem] * B[smem]^T, issued by ONE thread.
__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                         uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(0u)
      : "memory");
}

// mbarrier arrives when all MMAs issued so far by this thread have completed (implies
// tcgen05.fence::before_thread_sync).
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_addr(bar))
               : "memory");
}

// 32 lanes x 32 consecutive columns: thread t of the warp gets row (lane base + t).
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]),
        "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]),
        "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]),
        "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

}  // namespace umma
}  // namespace b200mc
