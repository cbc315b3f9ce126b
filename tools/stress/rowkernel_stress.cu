// Stand-alone reproducibility stress of gn_res_spk_ln (rowops.cuh): the kernel alone, fixed inputs, N launches, every
// output compared with the first launch's on the device; prints which rows ever differed.
//   nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -I serenade_b200/csrc tools/stress/rowkernel_stress.cu -o <bin>
//   <bin> [launches] [T] [rows_per_block]
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "rowops.cuh"

using namespace srn;
using OT = __nv_bfloat16;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ float hash01(unsigned long long i, unsigned s) {
  unsigned long long z = (i + 1) * 0x9E3779B97F4A7C15ull + s * 0xBF58476D1CE4E5B9ull;
  z ^= z >> 31; z *= 0x94D049BB133111EBull; z ^= z >> 29;
  return (float)(z & 0xffffff) / 16777216.0f;
}
__global__ void fill_bf16(OT* p, long long n, unsigned s, float scale) {
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n; i += gridDim.x * 256ll) p[i] = __float2bfloat16_rn((hash01(i, s) - 0.5f) * scale);
}
__global__ void fill_f32(float* p, long long n, unsigned s, float scale, float off) {
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n; i += gridDim.x * 256ll) p[i] = (hash01(i, s) - 0.5f) * scale + off;
}
// per row: does any of its 512 channels differ (hres f32 words / ln bf16 halves compared as raw bits)?
__global__ void compare_rows(const unsigned* a, const unsigned* b, const unsigned short* c, const unsigned short* d, long long rows,
                             unsigned* bad_rows) {
  const long long r = blockIdx.x;
  bool bad = false;
  for (int i = threadIdx.x; i < 512; i += blockDim.x) bad |= a[r * 512 + i] != b[r * 512 + i] || c[r * 512 + i] != d[r * 512 + i];
  if (__syncthreads_or(bad) && threadIdx.x == 0) atomicAdd(&bad_rows[r], 1u);
}

// Producer that rewrites dst (n_rows x 1024 B) from src before every consumer launch, like the GEMM epilogue does:
// mode 1 = generic 16-byte stores; mode 2 = rows staged in shared memory and written with bulk async stores
// (cp.async.bulk global <- shared, commit, wait_group 0 before exit: the async-proxy path of a TMA-store epilogue).
__global__ void __launch_bounds__(256) producer(const uint4* __restrict__ src, uint4* __restrict__ dst, long long n_rows, int mode) {
  __shared__ __align__(128) uint4 box[16 * 64];      // 16 rows x 1024 B
  for (long long r0 = blockIdx.x * 16ll; r0 < n_rows; r0 += gridDim.x * 16ll) {
    const int nr = (int)min(16ll, n_rows - r0);
    if (mode == 1) {
      for (int i = threadIdx.x; i < nr * 64; i += 256) dst[r0 * 64 + i] = src[r0 * 64 + i];
    } else {
      if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      __syncthreads();
      for (int i = threadIdx.x; i < nr * 64; i += 256) box[i] = src[r0 * 64 + i];
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();
      if (threadIdx.x == 0) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + r0 * 64), "r"(smem_u32(box)),
                     "r"(nr * 1024)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
  }
  if (mode == 2 && threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

int main(int argc, char** argv) {
  const int launches = argc > 1 ? atoi(argv[1]) : 20000;
  const int T = argc > 2 ? atoi(argv[2]) : 1000;
  const int rpb = argc > 3 ? atoi(argv[3]) : 56;
  const int mode = argc > 4 ? atoi(argv[4]) : 0;      // 0: static inputs; 1 / 2: rewritten before every launch (see producer)
  const int B = 16, mtiles = 2 * ((T + 255) / 256);
  const long long rows = (long long)B * T, n = rows * 512;
  OT *y, *res, *ln, *ln0, *ln1, *ysrc[2], *rsrc[2];
  float *part, *gamma, *beta, *sc, *sb, *lg, *lb, *hres, *hres0, *hres1;
  uint8_t* mask;
  unsigned* bad;
  CK(cudaMalloc(&y, n * 2)); CK(cudaMalloc(&res, n * 2)); CK(cudaMalloc(&ln, n * 2)); CK(cudaMalloc(&ln0, n * 2));
  CK(cudaMalloc(&hres, n * 4)); CK(cudaMalloc(&hres0, n * 4)); CK(cudaMalloc(&hres1, n * 4)); CK(cudaMalloc(&ln1, n * 2));
  for (int k = 0; k < 2; ++k) { CK(cudaMalloc(&ysrc[k], n * 2)); CK(cudaMalloc(&rsrc[k], n * 2)); }
  CK(cudaMalloc(&part, (size_t)B * mtiles * 16 * 4));
  CK(cudaMalloc(&gamma, 512 * 4)); CK(cudaMalloc(&beta, 512 * 4)); CK(cudaMalloc(&lg, 512 * 4)); CK(cudaMalloc(&lb, 512 * 4));
  CK(cudaMalloc(&sc, B * 512 * 4)); CK(cudaMalloc(&sb, B * 512 * 4));
  CK(cudaMalloc(&mask, rows)); CK(cudaMalloc(&bad, rows * 4));
  CK(cudaMemset(mask, 1, rows)); CK(cudaMemset(bad, 0, rows * 4));
  fill_bf16<<<1024, 256>>>(y, n, 1, 4.0f);
  fill_bf16<<<1024, 256>>>(res, n, 2, 2.0f);
  for (int k = 0; k < 2; ++k) { fill_bf16<<<1024, 256>>>(ysrc[k], n, 1 + 10 * k, 4.0f); fill_bf16<<<1024, 256>>>(rsrc[k], n, 2 + 10 * k, 2.0f); }
  fill_f32<<<8, 256>>>(gamma, 512, 3, 0.4f, 1.0f); fill_f32<<<8, 256>>>(beta, 512, 4, 0.2f, 0.f);
  fill_f32<<<8, 256>>>(lg, 512, 5, 0.4f, 1.0f); fill_f32<<<8, 256>>>(lb, 512, 6, 0.2f, 0.f);
  fill_f32<<<32, 256>>>(sc, B * 512, 7, 1.0f, 1.0f); fill_f32<<<32, 256>>>(sb, B * 512, 8, 0.5f, 0.f);
  {  // GroupNorm partial sums: per 128-row tile and group (sum ~ 0, sum of squares ~ rows * 64 * var(y))
    std::vector<float> h((size_t)B * mtiles * 16);
    for (size_t i = 0; i < h.size(); i += 2) { h[i] = 3.0f + (float)(i % 7); h[i + 1] = 128.0f * 64.0f * 1.4f; }
    CK(cudaMemcpy(part, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
  }
  dim3 grid((T + rpb - 1) / rpb, B);
  auto launch = [&](float* h, OT* l) {
    gn_res_spk_ln<OT><<<grid, 256>>>(y, part, mtiles, T, (long long)T * 512, gamma, beta, res, mask, T, 1, sc, sb,
                                                            lg, lb, h, l, rpb, nullptr);
  };
  auto produce = [&](int k) {
    producer<<<296, 256>>>((const uint4*)ysrc[k], (uint4*)y, rows, mode);
    producer<<<296, 256>>>((const uint4*)rsrc[k], (uint4*)res, rows, mode);
  };
  if (mode) { produce(1); CK(cudaDeviceSynchronize()); launch(hres1, ln1); CK(cudaDeviceSynchronize()); produce(0); }
  CK(cudaDeviceSynchronize());
  launch(hres0, ln0);
  CK(cudaDeviceSynchronize());
  for (int i = 0; i < launches; ++i) {
    const int k = mode ? (i & 1) : 0;
    if (mode) produce(k);
    launch(hres, ln);
    compare_rows<<<(unsigned)rows, 128>>>((const unsigned*)hres, (const unsigned*)(k ? hres1 : hres0), (const unsigned short*)ln,
                                          (const unsigned short*)(k ? ln1 : ln0), rows, bad);
  }
  CK(cudaDeviceSynchronize());
  std::vector<unsigned> hb(rows);
  CK(cudaMemcpy(hb.data(), bad, rows * 4, cudaMemcpyDeviceToHost));
  long long nbad = 0, events = 0;
  for (long long r = 0; r < rows; ++r) if (hb[r]) { ++nbad; events += hb[r]; }
  printf("mode %d: ", mode);
  printf("%d launches, T=%d rows_per_block=%d grid=(%u,%u): %lld rows ever differed (%lld row-events)\n", launches, T, rpb, grid.x, grid.y, nbad, events);
  int shown = 0;
  for (long long r = 0; r < rows && shown < 40; ++r)
    if (hb[r]) { printf("  b=%lld row=%lld (block %lld, row-in-block %lld, stage-iter %lld, warp %lld) x%u\n", r / T, r % T, (r % T) / rpb, (r % T) % rpb, ((r % T) % rpb) / 16, (((r % T) % rpb) % 16) / 2, hb[r]); ++shown; }
  return 0;
}
