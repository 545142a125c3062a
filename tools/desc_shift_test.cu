// Hardware probe: can a tcgen05 shared-memory matrix descriptor (K-major, SWIZZLE_128B) start at a row that
// is NOT a multiple of 8 (start address = base + s * 128 B), so that dilated conv taps become descriptor
// offsets into one resident activation strip?  Tries base_offset = 0 and base_offset = (addr >> 7) & 7.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I../speakerverification_b200/csrc -o desc_shift_test desc_shift_test.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "svb_ptx.cuh"
using namespace svb;

static constexpr int ROWS = 160, K = 64, N = 64;

__global__ void probe(const float* A, const float* W, float* D, int shift, int use_base_offset) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const uint32_t a_s = base, w_s = base + ROWS * 128;
  // fill A [ROWS x 64] and W [64 x 64] in the absolute-address swizzle (what TMA SWIZZLE_128B produces)
  for (int i = threadIdx.x; i < ROWS * K; i += blockDim.x) {
    const int r = i / K, c = i % K;
    const uint32_t addr = a_s + r * 128 + (((c >> 3) ^ (r & 7)) << 4) + (c & 7) * 2;
    __nv_bfloat16 v = __float2bfloat16(A[i]);
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"(*reinterpret_cast<unsigned short*>(&v)));
  }
  for (int i = threadIdx.x; i < N * K; i += blockDim.x) {
    const int r = i / K, c = i % K;
    const uint32_t addr = w_s + r * 128 + (((c >> 3) ^ (r & 7)) << 4) + (c & 7) * 2;
    __nv_bfloat16 v = __float2bfloat16(W[i]);
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"(*reinterpret_cast<unsigned short*>(&v)));
  }
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&slot), 64); tmem_relinquish(); }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x == 0) {
    const uint32_t a_addr = a_s + shift * 128;
    uint64_t ad = umma_desc_sw128(a_addr);
    if (use_base_offset) ad |= static_cast<uint64_t>((a_addr >> 7) & 7) << 49;
    const uint64_t bd = umma_desc_sw128(w_s);
    for (int k = 0; k < 4; ++k) umma_bf16(tm, ad + 2 * k, bd + 2 * k, umma_idesc_bf16(128, N), k ? 1u : 0u);
    umma_commit(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0);
  tc_fence_after();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int c = 0; c < N; c += 32) {
    uint32_t r[32];
    tmem_ld32(tm + c + (static_cast<uint32_t>(warp * 32) << 16), r);
    tmem_ld_wait();
    for (int i = 0; i < 32; ++i) D[(warp * 32 + lane) * N + c + i] = __uint_as_float(r[i]);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 64);
}

int main() {
  std::vector<float> A(ROWS * K), W(N * K), D(128 * N);
  srand(1);
  for (auto& v : A) v = static_cast<float>(rand() % 17 - 8);
  for (auto& v : W) v = static_cast<float>(rand() % 9 - 4);
  float *dA, *dW, *dD;
  cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dW, W.size() * 4); cudaMalloc(&dD, D.size() * 4);
  cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dW, W.data(), W.size() * 4, cudaMemcpyHostToDevice);
  const int smem = ROWS * 128 + N * 128 + 1024;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int ubo = 0; ubo < 2; ++ubo)
    for (int s = 0; s <= 12; ++s) {
      cudaMemset(dD, 0, D.size() * 4);
      probe<<<1, 128, smem>>>(dA, dW, dD, s, ubo);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("shift %d base_offset %d: CUDA error %s\n", s, ubo, cudaGetErrorString(e)); return 1; }
      cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
      int bad = 0;
      for (int m = 0; m < 128; ++m)
        for (int n = 0; n < N; ++n) {
          float ref = 0;
          for (int k = 0; k < K; ++k) ref += A[(m + s) * K + k] * W[n * K + k];
          if (ref != D[m * N + n]) ++bad;
        }
      printf("shift %2d base_offset_field %d: %s (%d mismatches)\n", s, ubo, bad ? "FAIL" : "ok", bad);
    }
  return 0;
}
