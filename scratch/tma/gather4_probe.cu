// build: nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o gather4_probe gather4_probe.cu ; run: ./gather4_probe 1
// probe: cp.async.bulk.tensor.2d tile::gather4 on sm_100a -- which box shape does the tensor map need, and how do the
// four rows land in shared memory?  Also the plain 2D tile copy of 64 contiguous rows x 16 doubles.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                             const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__global__ void k_probe(const __grid_constant__ CUtensorMap mapG, const __grid_constant__ CUtensorMap mapT, int col, int r0, int r1,
                        int r2, int r3, int trow, double *out) {
   __shared__ __align__(128) double buf[68][16];
   __shared__ __align__(8) unsigned long long bar;
   const int tid = threadIdx.x;
   for (int i = tid; i < 68 * 16; i += blockDim.x) (&buf[0][0])[i] = -1.0;
   if (tid == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
   }
   __syncthreads();
   if (tid == 0) {
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(4 * 128 + 64 * 128) : "memory");
      asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
                   ::"r"(smem_u32(&buf[0][0])), "l"(&mapG), "r"(col), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(smem_u32(&bar))
                   : "memory");
      asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                   ::"r"(smem_u32(&buf[4][0])), "l"(&mapT), "r"(col), "r"(trow), "r"(smem_u32(&bar))
                   : "memory");
   }
   unsigned done = 0;
   for (int spin = 0; spin < (1 << 24) && !done; ++spin)
      asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(smem_u32(&bar)), "r"(0) : "memory");
   if (!done && tid == 0) printf("TIMEOUT waiting for the copies\n");
   __syncthreads();
   for (int i = tid; i < 68 * 16; i += blockDim.x) out[i] = (&buf[0][0])[i];
}

int main(int argc, char **argv) {
   const int boxRowsG = argc > 1 ? atoi(argv[1]) : 1;
   const int R = 1000, K = 64;
   std::vector<double> h((size_t)R * K);
   for (int r = 0; r < R; ++r)
      for (int k = 0; k < K; ++k) h[(size_t)r * K + k] = r * 100.0 + k;
   double *d, *out;
   cudaMalloc(&d, h.size() * 8);
   cudaMalloc(&out, 68 * 16 * 8);
   cudaMemcpy(d, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
   EncodeFn enc = nullptr;
   cudaDriverEntryPointQueryResult qr;
   if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&enc, cudaEnableDefault, &qr) != cudaSuccess || !enc) {
      printf("no cuTensorMapEncodeTiled\n");
      return 1;
   }
   CUtensorMap mapG, mapT;
   cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)R}, strides[1] = {(cuuint64_t)K * 8};
   cuuint32_t boxG[2] = {16, (cuuint32_t)boxRowsG}, boxT[2] = {16, 64}, es[2] = {1, 1};
   // 8-byte elements: there is no FLOAT64 gather restriction known; try FLOAT64 first
   CUresult rc = enc(&mapG, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, d, dims, strides, boxG, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
   printf("encode gather map (box rows %d): %d\n", boxRowsG, (int)rc);
   rc = enc(&mapT, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, d, dims, strides, boxT, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
            CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
   printf("encode tile map: %d\n", (int)rc);
   k_probe<<<1, 128>>>(mapG, mapT, 32, 7, 500, 3, 999, 200, out);
   cudaError_t e = cudaDeviceSynchronize();
   printf("kernel: %s\n", cudaGetErrorString(e));
   std::vector<double> o(68 * 16);
   cudaMemcpy(o.data(), out, o.size() * 8, cudaMemcpyDeviceToHost);
   for (int r = 0; r < 6; ++r) printf("smem row %d: %.0f %.0f ... %.0f\n", r, o[r * 16], o[r * 16 + 1], o[r * 16 + 15]);
   printf("smem row 67: %.0f ... %.0f\n", o[67 * 16], o[67 * 16 + 15]);
   int ok = 1;
   const int rows[4] = {7, 500, 3, 999};
   for (int i = 0; i < 4; ++i)
      for (int k = 0; k < 16; ++k) ok &= o[i * 16 + k] == rows[i] * 100.0 + 32 + k;
   for (int i = 0; i < 64; ++i)
      for (int k = 0; k < 16; ++k) ok &= o[(4 + i) * 16 + k] == (200 + i) * 100.0 + 32 + k;
   printf(ok ? "PROBE_OK\n" : "PROBE_MISMATCH\n");
   return 0;
}
