// Probe: register layout of tcgen05.ld.16x256b.x8 (which TMEM lane / column lands in which thread / register).
// TMEM is filled through tcgen05.st.32x32b (thread = lane, register = column) with value = lane * 1000 + column.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o scripts/probe/tmem_ld_shapes scripts/probe/tmem_ld_shapes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void probe(uint32_t* out) {
  __shared__ uint32_t slot;
  const int lane = threadIdx.x;
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(64u) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  __syncwarp();
  const uint32_t t = slot;
  uint32_t v[32];
  for (int half = 0; half < 2; ++half) {
    for (int c = 0; c < 32; ++c) v[c] = lane * 1000 + half * 32 + c;
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
        ::"r"(t + half * 32), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
        "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]),
        "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
        : "memory");
  }
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  __syncwarp();
  for (int h16 = 0; h16 < 2; ++h16) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
          "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
          "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(t + ((uint32_t)(h16 * 16) << 16)));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 32; ++i) out[(h16 * 32 + lane) * 32 + i] = r[i];
  }
  __syncwarp();
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(t), "r"(64u) : "memory");
}

int main() {
  uint32_t* d;
  cudaMalloc(&d, 2 * 32 * 32 * 4);
  probe<<<1, 32>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
  static uint32_t h[2 * 32 * 32];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  for (int h16 = 0; h16 < 2; ++h16)
    for (int lane = 0; lane < 8; ++lane) {
      printf("half %d thread %2d:", h16, lane);
      for (int i = 0; i < 12; ++i) printf(" r%d=(%u,%u)", i, h[(h16 * 32 + lane) * 32 + i] / 1000, h[(h16 * 32 + lane) * 32 + i] % 1000);
      printf("\n");
    }
  // check the hypothesis: r[4k + 2a + b] of thread t = (lane h16*16 + t/4 + 8a, column 8k + 2(t%4) + b)
  int bad = 0;
  for (int h16 = 0; h16 < 2; ++h16)
    for (int t = 0; t < 32; ++t)
      for (int k = 0; k < 8; ++k)
        for (int a = 0; a < 2; ++a)
          for (int b = 0; b < 2; ++b) {
            const uint32_t got = h[(h16 * 32 + t) * 32 + 4 * k + 2 * a + b];
            const uint32_t want = (h16 * 16 + t / 4 + 8 * a) * 1000 + 8 * k + 2 * (t % 4) + b;
            bad += got != want;
          }
  printf("hypothesis r[4k+2a+b] = (lane 16h + t/4 + 8a, col 8k + 2(t%%4) + b): %s (%d mismatches)\n", bad ? "WRONG" : "confirmed", bad);
  return 0;
}
