// Probe: does cuTensorMapEncodeTiled accept a row stride SMALLER than the row extent (overlapping
// rows: a Toeplitz / sliding-window view of a 1-D signal), and does the TMA load return
// x[r * stride + c]?  Used by the sinc-conv front-end of RawNet2 (rawnet_gemm.cu).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o tma_overlap_probe tma_overlap_probe.cu -lcuda
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <vector>

__global__ void probe(const __grid_constant__ CUtensorMap tm, int c0, int r0, __nv_bfloat16* out) {
  __shared__ __align__(1024) __nv_bfloat16 tile[8 * 64];
  __shared__ __align__(8) uint64_t bar;
  const uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar);
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(tile);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b));
    asm volatile("fence.mbarrier_init.release.cluster;");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(8 * 64 * 2));
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(s), "l"((uint64_t)&tm), "r"(b), "r"(c0), "r"(r0) : "memory");
    uint32_t done = 0;
    while (!done)
      asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.b32 %0,1,0,p;\n}" : "=r"(done) : "r"(b));
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 8 * 64; i += blockDim.x) out[i] = tile[i];
}

int main() {
  const int n = 1 << 16, stride = 48;
  std::vector<__nv_bfloat16> h(n);
  for (int i = 0; i < n; ++i) h[i] = __float2bfloat16((float)(i % 251));
  __nv_bfloat16 *d, *o;
  cudaMalloc(&d, n * 2); cudaMalloc(&o, 8 * 64 * 2);
  cudaMemcpy(d, h.data(), n * 2, cudaMemcpyHostToDevice);
  CUtensorMap tm;
  cuuint64_t dims[2] = {(cuuint64_t)20000, (cuuint64_t)((n - 20000) / stride)};
  cuuint64_t strides[1] = {(cuuint64_t)stride * 2};
  cuuint32_t box[2] = {64, 8}, es[2] = {1, 1};
  cuInit(0);
  CUresult r = cuTensorMapEncodeTiled(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, dims, strides, box, es,
                                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                      CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode (row stride %d elements < row extent %d): CUresult %d\n", stride, 20000, (int)r);
  if (r != CUDA_SUCCESS) return 1;
  int bad = 0;
  const int tests[3][2] = {{0, 0}, {256, 5}, {10048, 100}};
  for (auto& t : tests) {
    probe<<<1, 128>>>(tm, t[0], t[1], o);
    std::vector<__nv_bfloat16> g(8 * 64);
    cudaError_t e = cudaMemcpy(g.data(), o, 8 * 64 * 2, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(e)); return 1; }
    for (int rr = 0; rr < 8; ++rr)
      for (int c = 0; c < 64; ++c) {
        const float want = (float)(((t[1] + rr) * stride + t[0] + c) % 251);
        if (__bfloat162float(g[rr * 64 + c]) != want) ++bad;
      }
    printf("box at (col %d, row %d): mismatches so far %d\n", t[0], t[1], bad);
  }
  printf(bad ? "FAIL\n" : "PASS: overlapping-row TMA view works\n");
  return bad != 0;
}
