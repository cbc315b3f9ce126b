// GEMM epilogue: per-warp staging of 32-row output boxes in shared memory + TMA bulk stores.
//
// Each epilogue thread owns one accumulator row.  Writing that row straight to global memory makes every
// warp-wide store touch 32 different cache lines (ncu / experiments showed the K=512 GEMMs bound by exactly
// that), so rows are first written to a 128-byte-swizzled smem box (conflict-free 16-byte stores) and one
// lane issues a TMA store of the 32-row x 128-byte box, which writes full lines and clips rows >= M and
// columns >= N in hardware.
#pragma once

namespace srn {

// Per-warp staging area: two 4 KB boxes (double buffered against the asynchronous TMA reads).
struct Stager {
  uint32_t base;     // shared address of this warp's 8 KB, 1024-byte aligned
  uint32_t buf;      // which box is being filled
  uint32_t fill;     // bytes of the current 128-byte row already written
  uint32_t one;      // 1: a single 4 KB box (16 epilogue warps) - every fill waits for the previous store's reads
  __device__ __forceinline__ uint32_t box() const { return base + buf * 4096; }
};

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}

// Before the first write into a box: the TMA store issued two boxes ago must have finished reading it.
__device__ __forceinline__ void stager_acquire(Stager& s, int lane) {
  if (s.fill == 0) {
    if (lane == 0) {
      if (s.one) tma_store_wait_read<0>();
      else tma_store_wait_read<1>();
    }
    __syncwarp();
  }
}

// Append N consecutive output values of this lane's row (row index inside the box = lane), SWIZZLE_128B.
template <typename OutT, int N>
__device__ __forceinline__ void stager_put(Stager& s, int lane, const float (&v)[N]) {
  constexpr int kBytes = N * (int)sizeof(OutT);
  static_assert(kBytes % 16 == 0 && kBytes <= 128, "row pieces are whole 16-byte chunks");
  const uint32_t row = s.box() + lane * 128;
  const uint32_t c0 = s.fill >> 4;
#pragma unroll
  for (int k = 0; k < kBytes / 16; ++k) {
    const uint32_t addr = row + (((c0 + k) ^ (lane & 7)) << 4);
    if constexpr (sizeof(OutT) == 4) {
      st_shared_v4(addr, __float_as_uint(v[4 * k]), __float_as_uint(v[4 * k + 1]), __float_as_uint(v[4 * k + 2]),
                   __float_as_uint(v[4 * k + 3]));
    } else {
      st_shared_v4(addr, pack_bf16x2(v[8 * k], v[8 * k + 1]), pack_bf16x2(v[8 * k + 2], v[8 * k + 3]),
                   pack_bf16x2(v[8 * k + 4], v[8 * k + 5]), pack_bf16x2(v[8 * k + 6], v[8 * k + 7]));
    }
  }
  s.fill += kBytes;
}

// When the 128-byte rows are complete: store the box.  col = first output column (elements) of the box,
// row0 = first row, batch = batch coordinate.
// The writes are made visible to the async proxy and one lane issues the TMA bulk store of the 32 x 128 B box (a direct
// st.global epilogue was measured in round 1 and is 2 % slower: the bytes, not the store path, are what costs).
template <typename OutT>
__device__ __forceinline__ void stager_flush(Stager& s, int lane, const GemmParams& p, int col, int row0, int batch,
                                             bool reduce_add = false, const CUtensorMap* map = nullptr) {
  if (s.fill < 128) return;
  if (map == nullptr) map = &p.mapD;
  fence_proxy_async();
  __syncwarp();
  if (lane == 0) {
    if (reduce_add) {
      // out += box, the addition done by the L2 (f32, every element touched once: deterministic): the residual
      // stream is updated in place without the SM ever reading it
      asm volatile("cp.reduce.async.bulk.tensor.3d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                       reinterpret_cast<uint64_t>(map)),
                   "r"(s.box()), "r"(col), "r"(row0), "r"(batch)
                   : "memory");
    } else {
      asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                       reinterpret_cast<uint64_t>(map)),
                   "r"(s.box()), "r"(col), "r"(row0), "r"(batch)
                   : "memory");
    }
    tma_store_commit();
  }
  if (!s.one) s.buf ^= 1;
  s.fill = 0;
}

// V^T box for EPI_QKV: [32 d rows][32 s] with s contiguous, no swizzle; lane = s, v[i] = column d0 + i.
template <typename OutT>
__device__ __forceinline__ void stager_store_transposed(Stager& s, int lane, const float (&v)[32],
                                                        const CUtensorMap* map, int s0, int d0, int batch) {
  if (lane == 0) {
    if (s.one) tma_store_wait_read<0>();
    else tma_store_wait_read<1>();
  }
  __syncwarp();
  const uint32_t box = s.box();
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    const uint32_t addr = box + (i * 32 + lane) * (uint32_t)sizeof(OutT);
    if constexpr (sizeof(OutT) == 4) {
      asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(__float_as_uint(v[i])) : "memory");
    } else {
      const __nv_bfloat16 h = __float2bfloat16_rn(v[i]);
      asm volatile("st.shared.b16 [%0], %1;" ::"r"(addr), "h"(*reinterpret_cast<const uint16_t*>(&h)) : "memory");
    }
  }
  fence_proxy_async();
  __syncwarp();
  if (lane == 0) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                     reinterpret_cast<uint64_t>(map)),
                 "r"(box), "r"(s0), "r"(d0), "r"(batch)
                 : "memory");
    tma_store_commit();
  }
  if (!s.one) s.buf ^= 1;
  s.fill = 0;
}

}  // namespace srn
