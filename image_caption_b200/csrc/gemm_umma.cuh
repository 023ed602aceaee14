// gemm_umma.cuh - tcgen05 GEMM, C[M,N] = A[M,K] * W[N,K]^T with bf16 operands (kind::f16) or fp32 operands read as
// TF32 (kind::tf32, the fp32-class throughput mode CAPDEC_PREC_TF32) and fp32 accumulation in TMEM.  Both operands are K-major (A = activations, W = nn.Linear weight),
// staged by TMA into 128B-swizzled shared-memory tiles.
//
// One 128 x BLOCK_N output tile per CTA; 6 warps:
//   warp 0   TMA producer (one elected lane)
//   warp 1   TMEM allocator + MMA issuer (one elected lane issues tcgen05.mma)
//   warp 2-5 epilogue: tcgen05.ld -> registers -> Epi::store (each thread owns one row)
// Pipelines: full[stage]/empty[stage] (TMA <-> MMA) and tmem_full (MMA -> epilogue).
#pragma once
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace capdec {

constexpr int kUmmaBM = 128;
constexpr int kUmmaBK = 64;   // 64 bf16 = 128 B = one swizzle atom row
constexpr int kUmmaK = 16;    // K per tcgen05.mma for 16-bit inputs
constexpr int kUmmaThreads = 192;

__host__ __device__ constexpr int umma_stage_bytes(int block_n) { return (kUmmaBM + block_n) * kUmmaBK * 2; }
__host__ __device__ constexpr int umma_smem_bytes(int block_n, int stages) {
  return stages * umma_stage_bytes(block_n) + 1024 /*alignment slack*/ + 256 /*barriers*/;
}

// In = operand element type in global / shared memory: bf16 (64 per 128-byte swizzle row) or float (32 per row)
template <int BLOCK_N, typename Epi, typename In = bf16>
__global__ void __launch_bounds__(kUmmaThreads, 2)
gemm_umma_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_w,
                 int M, int N, int K, int stages, Epi epi, const DecState* st) {
  static_assert(BLOCK_N == 32 || BLOCK_N == 64 || BLOCK_N == 128 || BLOCK_N == 256, "BLOCK_N");
  if (st != nullptr && st->done) return;  // uniform: taken by every thread or none

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  constexpr int kABytes = kUmmaBM * kUmmaBK * 2;
  constexpr int kBBytes = BLOCK_N * kUmmaBK * 2;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + (size_t)stages * kABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + (size_t)stages * kBBytes);
  uint64_t* full = bars;               // [stages]
  uint64_t* empty = bars + stages;     // [stages]
  uint64_t* tmem_full = bars + 2 * stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * stages + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * kUmmaBM, n0 = blockIdx.x * BLOCK_N;
  constexpr int kBKe = 128 / (int)sizeof(In);  // elements per k-block (one 128-byte swizzle row)
  const int num_kb = K / kBKe;

  // TMA / tcgen05 instructions are issued by the converged warp under elect.sync (see gemm_swap.cuh: under
  // `if (lane == 0)` every instruction is wrapped in a uniformity loop and costs ~150 cycles)
  if (warp == 0) {
    if (ptx::elect_one()) {
      ptx::prefetch_tensormap(&tm_a);
      ptx::prefetch_tensormap(&tm_w);
      for (int s = 0; s < stages; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
      ptx::mbar_init(tmem_full, 1);
      ptx::fence_barrier_init();
    }
    __syncwarp();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, BLOCK_N);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    for (int kb = 0; kb < num_kb; ++kb) {
      const int s = kb % stages;
      const uint32_t ph = (kb / stages) & 1;
      ptx::mbar_wait(&empty[s], ph ^ 1);
      if (ptx::elect_one()) {
        ptx::mbar_expect_tx(&full[s], kABytes + kBBytes);
        ptx::tma_load_2d(smem_a + (size_t)s * kABytes, &tm_a, &full[s], kb * kBKe, m0);
        ptx::tma_load_2d(smem_b + (size_t)s * kBBytes, &tm_w, &full[s], kb * kBKe, n0);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    constexpr bool kTf32 = sizeof(In) == 4;
    constexpr uint32_t idesc = kTf32 ? ptx::umma_idesc_tf32(BLOCK_N) : ptx::umma_idesc_bf16(BLOCK_N);
    for (int kb = 0; kb < num_kb; ++kb) {
      const int s = kb % stages;
      const uint32_t ph = (kb / stages) & 1;
      ptx::mbar_wait(&full[s], ph);
      ptx::tc_fence_after();
      if (ptx::elect_one()) {
        const uint32_t a_addr = ptx::smem_u32(smem_a + (size_t)s * kABytes);
        const uint32_t b_addr = ptx::smem_u32(smem_b + (size_t)s * kBBytes);
#pragma unroll
        for (int k = 0; k < kUmmaBK / kUmmaK; ++k) {
          // advance inside the 128 B swizzle atom: 16 bf16 = 8 tf32 = 32 B per MMA
          if constexpr (kTf32)
            ptx::umma_tf32(tmem_base, ptx::umma_desc_sw128(a_addr + k * 32), ptx::umma_desc_sw128(b_addr + k * 32), idesc,
                           (kb | k) != 0 ? 1u : 0u);
          else
            ptx::umma_f16(tmem_base, ptx::umma_desc_sw128(a_addr + k * kUmmaK * 2),
                          ptx::umma_desc_sw128(b_addr + k * kUmmaK * 2), idesc, (kb | k) != 0 ? 1u : 0u);
        }
        ptx::umma_commit(&empty[s]);                        // smem slot reusable once these MMAs retire
        if (kb == num_kb - 1) ptx::umma_commit(tmem_full);  // accumulator complete (same thread as its MMAs)
      }
      __syncwarp();
    }
  } else {
    // epilogue warps 2..5: TMEM lane quarter = warp % 4
    const int quarter = warp & 3;
    const int row = m0 + quarter * 32 + lane;
    ptx::mbar_wait(tmem_full, 0);
    ptx::tc_fence_after();
#pragma unroll 1
    for (int c = 0; c < BLOCK_N; c += 32) {
      float v[32];
      ptx::tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)c, v);
      if (row < M && n0 + c < N) epi.template store<32>(row, n0 + c, v);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, BLOCK_N);
}

}  // namespace capdec
