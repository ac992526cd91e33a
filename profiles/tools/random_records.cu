// random_records.cu — what HBM delivers for reads of RANDOMLY placed records of R bytes (16-byte aligned),
// the access pattern of a drift check over small records whose shard copies are not in the controller's
// order.  One warp reads `UNROLL` records per step (each lane one 16-byte vector of each record for R <= 512,
// looping for longer ones), all loads of a step issued before any is used.  Offsets come from a hash of the
// record number: no index traffic.  Prints GB/s per record size: the roofline small-record scans are held to.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o random_records random_records.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t mix(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

template <int UNROLL>
__global__ void __launch_bounds__(256) k_read(const uint8_t* __restrict__ buf, uint64_t slots, uint32_t rec_bytes, uint64_t nrec,
                                              uint32_t* __restrict__ sink, int sequential) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  const uint32_t nvec = rec_bytes >> 4;
  const uint32_t lanes_per_rec = nvec < 32 ? nvec : 32;  // short records share a warp
  const uint32_t rec_per_warp = 32 / lanes_per_rec, sub = lane / lanes_per_rec, l = lane % lanes_per_rec;
  uint32_t acc = 0;
  for (uint64_t r0 = warp * rec_per_warp * UNROLL; r0 < nrec; r0 += nwarps * rec_per_warp * UNROLL) {
    const uint4* p[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const uint64_t r = r0 + (uint64_t)u * rec_per_warp + sub;
      const uint64_t slot = (sequential ? r : mix(r)) & (slots - 1);  // slots is a power of two
      p[u] = reinterpret_cast<const uint4*>(buf + slot * rec_bytes);
    }
    for (uint32_t v = l; v < nvec; v += lanes_per_rec) {
      uint4 x[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) x[u] = ldg_stream(p[u] + v);
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) acc |= x[u].x ^ x[u].y ^ x[u].z ^ x[u].w;
    }
  }
  if (acc == 0x12345678u) sink[0] = acc;
}

int main() {
  const uint64_t bytes = 16ull << 30;
  uint8_t* buf;
  uint32_t* sink;
  cudaMalloc(&buf, bytes);
  cudaMalloc(&sink, 4);
  cudaMemset(buf, 1, bytes);
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  printf("{\"sms\": %d, \"buffer_gb\": %.1f, \"rows\": [\n", sms, bytes / 1e9);
  const uint32_t sizes[] = {32, 64, 128, 256, 512, 1024, 2048, 4096, 16384};
  bool first = true;
  for (uint32_t R : sizes)
    for (int seq = 0; seq < 2; ++seq)
      for (int ctas = 4; ctas <= 8; ctas += 4) {
        const uint64_t slots = bytes / R;
        const uint64_t nrec = (8ull << 30) / R;  // 8 GB read per launch
        float best = 1e9f;
        for (int it = 0; it < 4; ++it) {
          cudaEventRecord(a);
          k_read<8><<<sms * ctas, 256>>>(buf, slots, R, nrec, sink, seq);
          cudaEventRecord(b);
          cudaEventSynchronize(b);
          float ms;
          cudaEventElapsedTime(&ms, a, b);
          if (it > 0 && ms < best) best = ms;
        }
        printf("%s{\"record_bytes\": %u, \"order\": \"%s\", \"ctas_per_sm\": %d, \"ms\": %.3f, \"gb_per_s\": %.0f}", first ? "" : ",\n", R,
               seq ? "sequential" : "random", ctas, best, nrec * R / (best * 1e6));
        first = false;
      }
  printf("\n]}\n");
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    fprintf(stderr, "cuda: %s\n", cudaGetErrorString(e));
    return 1;
  }
  return 0;
}
