// Experiments that pin two hardware facts the halo-resident convolution kernel depends on (run on a B200):
//  (1) tcgen05.mma SS-mode issue rate for 128 x N x 16 bf16, N in {64,128,192,256} (is small N smem-read bound?)
//  (2) SWIZZLE_128B K-major A descriptors whose start address is shifted by whole 128-byte rows (not 1024-aligned)
//      and whose 8-row groups sit SBO = 14*128 B apart: does the MMA read rows start+g*SBO+r*128 with the swizzle
//      anchored on absolute smem address bits (what TMA wrote), with base_offset = 0 or (start>>7)&7 ?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I raygun_b200/csrc scripts/exp_umma.cu -o gpurun_out/exp_umma
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "sm100_ptx.cuh"
using namespace rg;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode() {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  return (EncodeTiledFn)p;
}

// ---------------- (1) issue rate ----------------
template <int N, int MODE>   // MODE 0: rotate A/B addresses through a 96 KB window; 1: same A and B every time
__global__ void __launch_bounds__(128, 1) rate_kernel(int iters, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem_raw)[i] = 0;
  if (threadIdx.x == 0) { ptx::mbar_init(ptx::smem_u32(&bar), 1); ptx::fence_mbar_init(); }
  if (threadIdx.x < 32) ptx::tmem_alloc<512>(ptx::smem_u32(&slot));
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = ptx::idesc_bf16(128, N, 0, 0);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      const uint32_t a = base + (MODE ? 0 : (it & 3) * 16384);
      const uint32_t b = base + 65536 + (MODE ? 0 : (it & 1) * 32768);
      const uint64_t ad = ptx::smem_desc_sw128(a, 16, 1024), bd = ptx::smem_desc_sw128(b, 16, 1024);
#pragma unroll
      for (int k = 0; k < 4; ++k) ptx::umma_bf16(tmem + (uint32_t)((it & 1) * 256), ad + 2 * k, bd + 2 * k, idesc, 1u);
    }
    ptx::umma_commit(ptx::smem_u32(&bar));
    ptx::mbar_wait(ptx::smem_u32(&bar), 0);
    long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc<512>(tmem); }
}

template <int N, int MODE> static void run_rate(const char* name, long long* d_out) {
  const int iters = 4000;
  cudaFuncSetAttribute(rate_kernel<N, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 161 * 1024);
  rate_kernel<N, MODE><<<148, 128, 161 * 1024>>>(iters, d_out);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < 148; ++i) avg += (double)h[i];
  avg /= 148.0 * iters * 4;
  printf("rate %-28s N=%3d: %.2f clk per 128xNx16 MMA (ideal %.1f) -> %.1f%% of tensor peak  [%s]\n", name, N, avg, N / 2.0,
         100.0 * (N / 2.0) / avg, cudaGetErrorString(e));
}

// ---------------- (2) shifted descriptors ----------------
constexpr int SRC_ROWS = 320;   // 308 used
__global__ void __launch_bounds__(128, 1) shift_kernel(const __grid_constant__ CUtensorMap src_map,
                                                       const __grid_constant__ CUtensorMap id_map, int shift_rows, int sbo_rows,
                                                       int use_base_offset, int mn_major, float* out /*128x64*/) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bars[2];
  __shared__ uint32_t slot;
  const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t smem_a = base, smem_b = base + 64 * 1024;
  if (threadIdx.x == 0) {
    ptx::mbar_init(ptx::smem_u32(&bars[0]), 1); ptx::mbar_init(ptx::smem_u32(&bars[1]), 1); ptx::fence_mbar_init();
  }
  if (threadIdx.x < 32) ptx::tmem_alloc<64>(ptx::smem_u32(&slot));
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    const uint32_t fb = ptx::smem_u32(&bars[0]);
    ptx::mbar_arrive_expect_tx(fb, 2 * 256 * 128 + 64 * 128);   // OOB rows of the 2nd box are zero-filled but counted
    ptx::tma_load_2d(smem_a, &src_map, fb, 0, 0);                    // rows 0..255
    ptx::tma_load_2d(smem_a + 256 * 128, &src_map, fb, 0, 256);      // rows 256..319
    ptx::tma_load_2d(smem_b, &id_map, fb, 0, 0);
    ptx::mbar_wait(fb, 0);
    ptx::tc_fence_after();
    const uint32_t start = smem_a + shift_rows * 128;
    uint64_t ad, bd;
    uint32_t idesc;
    if (!mn_major) {
      ad = ptx::smem_desc_sw128(start, 16, sbo_rows * 128);
      bd = ptx::smem_desc_sw128(smem_b, 16, 1024);
      idesc = ptx::idesc_bf16(128, 64, 0, 0);
    } else {
      // A is MN-major: smem rows are K (here: the 64 src rows of 8 groups), columns (64) are M ... handled by host expectations
      ad = ptx::smem_desc_sw128(start, 16384, sbo_rows * 128);
      bd = ptx::smem_desc_sw128(smem_b, 16, 1024);
      idesc = ptx::idesc_bf16(128, 64, 1, 0);   // rows 64..127 read the slab at +LBO (ignored by the check)
    }
    if (use_base_offset) ad |= (uint64_t)((start >> 7) & 7) << 49;
    if (!mn_major) {
      for (int k = 0; k < 4; ++k) ptx::umma_bf16(tmem, ad + 2 * k, bd + 2 * k, idesc, k ? 1u : 0u);
    } else {
      // K = 64 src rows (identity B picks k -> n), 4 MMAs of K=16 = two 8-row groups each
      for (int k = 0; k < 4; ++k) {
        uint64_t adk = ptx::smem_desc_sw128(start + k * 2 * sbo_rows * 128, 16384, sbo_rows * 128);
        if (use_base_offset) adk |= (uint64_t)(((start + k * 2 * sbo_rows * 128) >> 7) & 7) << 49;
        ptx::umma_bf16(tmem, adk, bd + 2 * k, idesc, k ? 1u : 0u);
      }
    }
    ptx::umma_commit(ptx::smem_u32(&bars[1]));
  }
  __syncthreads();
  ptx::mbar_wait(ptx::smem_u32(&bars[1]), 0);
  ptx::tc_fence_after();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t v[64];
  ptx::tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16), v);
  ptx::tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + 32, v + 32);
  ptx::tmem_ld_wait();
  for (int c = 0; c < 64; ++c) out[(warp * 32 + lane) * 64 + c] = __uint_as_float(v[c]);
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc<64>(tmem); }
}

int main() {
  long long* d_out;
  cudaMalloc(&d_out, 148 * sizeof(long long));
  run_rate<64, 0>("SS rotating operands", d_out);
  run_rate<128, 0>("SS rotating operands", d_out);
  run_rate<192, 0>("SS rotating operands", d_out);
  run_rate<256, 0>("SS rotating operands", d_out);
  run_rate<64, 1>("SS same operands", d_out);
  run_rate<128, 1>("SS same operands", d_out);
  run_rate<256, 1>("SS same operands", d_out);

  // ---- shifted descriptor test ----
  EncodeTiledFn enc = get_encode();
  std::vector<__nv_bfloat16> src(SRC_ROWS * 64), ident(64 * 64);
  for (int r = 0; r < SRC_ROWS; ++r)
    for (int c = 0; c < 64; ++c) src[r * 64 + c] = __float2bfloat16((float)(r * 64 + c) * (1.0f / 64.0f) - 100.0f);   // exact in bf16? use small ints
  for (int r = 0; r < SRC_ROWS; ++r)
    for (int c = 0; c < 64; ++c) src[r * 64 + c] = __float2bfloat16((float)((r * 7 + c * 3) % 251 - 125));
  for (int r = 0; r < 64; ++r)
    for (int c = 0; c < 64; ++c) ident[r * 64 + c] = __float2bfloat16(r == c ? 1.0f : 0.0f);
  __nv_bfloat16 *d_src, *d_id;
  float* d_res;
  cudaMalloc(&d_src, src.size() * 2); cudaMalloc(&d_id, ident.size() * 2); cudaMalloc(&d_res, 128 * 64 * 4);
  cudaMemcpy(d_src, src.data(), src.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(d_id, ident.data(), ident.size() * 2, cudaMemcpyHostToDevice);
  CUtensorMap src_map, id_map;
  {
    cuuint64_t dims[2] = {64, SRC_ROWS}; cuuint64_t gs[1] = {128}; cuuint32_t bx[2] = {64, 256}; cuuint32_t es[2] = {1, 1};
    CUresult r = enc(&src_map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d_src, dims, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    cuuint64_t dims2[2] = {64, 64}; cuuint32_t bx2[2] = {64, 64};
    CUresult r2 = enc(&id_map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d_id, dims2, gs, bx2, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r || r2) { printf("encode failed %d %d\n", (int)r, (int)r2); return 1; }
  }
  cudaFuncSetAttribute(shift_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
  std::vector<float> res(128 * 64);
  for (int mn = 0; mn < 2; ++mn)
    for (int ubo = 0; ubo < 2; ++ubo)
      for (int sbo_rows : {8, 14})
        for (int shift : {0, 1, 3, 7, 8, 14, 17, 14 * 6 + 6}) {
          cudaMemset(d_res, 0, 128 * 64 * 4);
          shift_kernel<<<1, 128, 80 * 1024>>>(src_map, id_map, shift, sbo_rows, ubo, mn, d_res);
          cudaError_t e = cudaDeviceSynchronize();
          if (e != cudaSuccess) { printf("shift kernel failed: %s\n", cudaGetErrorString(e)); return 1; }
          cudaMemcpy(res.data(), d_res, res.size() * 4, cudaMemcpyDeviceToHost);
          int bad = 0, first = -1;
          if (!mn) {
            for (int row = 0; row < 128; ++row)
              for (int c = 0; c < 64; ++c) {
                const int sr = shift + (row / 8) * sbo_rows + (row % 8);
                const float exp = __bfloat162float(src[sr * 64 + c]);
                if (res[row * 64 + c] != exp) { if (first < 0) first = row * 64 + c; ++bad; }
              }
          } else {
            // D[m][n] = sum_k A[m][k] I[n][k] = A[m][n]; A[m][k] = src[row(k)][m], row(k) = shift + (k/8)*sbo + k%8, M = 64
            for (int m = 0; m < 64; ++m)
              for (int n = 0; n < 64; ++n) {
                const int sr = shift + (n / 8) * sbo_rows + (n % 8);
                const float exp = __bfloat162float(src[sr * 64 + m]);
                if (res[m * 64 + n] != exp) { if (first < 0) first = m * 64 + n; ++bad; }
              }
          }
          printf("shift %s base_offset=%d sbo_rows=%2d shift=%3d : %s (%d mismatches, first at %d)\n", mn ? "MN-major" : "K-major ",
                 ubo, sbo_rows, shift, bad ? "MISMATCH" : "ok", bad, first);
        }
  return 0;
}
