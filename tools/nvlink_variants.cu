// nvlink_variants.cu -- developer microbenchmark (not part of the library), 2 GPUs, one process:
// does the way a kernel issues NVLink traffic change the achievable rate?
//   pull  : GPU0 reads 102 MB from GPU1   (a) LDG.256 per thread   (b) TMA bulk copies into smem
//   push  : GPU0 writes 102 MB to GPU1    (c) STG.256 per thread   (d) TMA bulk copies from smem
//   both  : pull and push in the same kernel / concurrently on two streams
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o nvlink_variants nvlink_variants.cu
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

#define CK(x)                                                                        \
  do {                                                                               \
    cudaError_t e = (x);                                                             \
    if (e != cudaSuccess) {                                                          \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
      exit(1);                                                                       \
    }                                                                                \
  } while (0)

struct f8 { float a[8]; };
__device__ __forceinline__ f8 ld8(const float* p) {
  f8 r;
  asm volatile("ld.global.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r.a[0]), "=f"(r.a[1]), "=f"(r.a[2]), "=f"(r.a[3]), "=f"(r.a[4]), "=f"(r.a[5]), "=f"(r.a[6]), "=f"(r.a[7])
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st8(float* p, const f8& r) {
  asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(r.a[0]), "f"(r.a[1]), "f"(r.a[2]),
               "f"(r.a[3]), "f"(r.a[4]), "f"(r.a[5]), "f"(r.a[6]), "f"(r.a[7])
               : "memory");
}

// (a) pull with LDG.256: remote -> local copy (persistent grid)
__global__ void __launch_bounds__(512) pull_ldg(const float* remote, float* local, long nvec8) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long q = (long)blockIdx.x * blockDim.x + threadIdx.x; q < nvec8; q += stride) st8(local + q * 8, ld8(remote + q * 8));
}
// (c) push with STG.256: local -> remote copy
__global__ void __launch_bounds__(512) push_stg(const float* local, float* remote, long nvec8) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long q = (long)blockIdx.x * blockDim.x + threadIdx.x; q < nvec8; q += stride) st8(remote + q * 8, ld8(local + q * 8));
}
// pull + push fused per thread (what the step kernel does)
__global__ void __launch_bounds__(512) both_ldst(const float* remote_src, float* local_dst, const float* local_src,
                                                 float* remote_dst, long nvec8) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long q = (long)blockIdx.x * blockDim.x + threadIdx.x; q < nvec8; q += stride) {
    f8 a = ld8(remote_src + q * 8), b = ld8(local_src + q * 8);
    st8(local_dst + q * 8, a);
    st8(remote_dst + q * 8, b);
  }
}

// one kernel, specialised blocks: even blocks pull, odd blocks push
__global__ void __launch_bounds__(512) split_ldst(const float* remote_src, float* local_dst, const float* local_src,
                                                  float* remote_dst, long nvec8) {
  const long stride = (long)(gridDim.x / 2) * blockDim.x;
  const long q0 = (long)(blockIdx.x / 2) * blockDim.x + threadIdx.x;
  if (blockIdx.x & 1)
    for (long q = q0; q < nvec8; q += stride) st8(remote_dst + q * 8, ld8(local_src + q * 8));
  else
    for (long q = q0; q < nvec8; q += stride) st8(local_dst + q * 8, ld8(remote_src + q * 8));
}

// TMA (cp.async.bulk) helpers
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  unsigned done = 0;
  while (!done)
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(done)
                 : "r"(smem_u32(bar)), "r"(parity)
                 : "memory");
}

// (b)/(d) TMA copy src -> dst through shared memory: one elected thread per CTA drives a ring of
// STAGES buffers of CHUNK bytes (load chunk i+STAGES-1 while storing chunk i).  Either side may be remote.
template <int CHUNK, int STAGES>
__global__ void __launch_bounds__(32) copy_tma(const char* src, char* dst, long nchunks) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long bar[STAGES];
  if (threadIdx.x != 0) return;
  for (int s = 0; s < STAGES; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[s])));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  // chunks handled by this CTA: blockIdx.x, blockIdx.x + gridDim.x, ...
  long mine = (nchunks - blockIdx.x + gridDim.x - 1) / gridDim.x;
  if (mine <= 0) return;
  auto issue = [&](long k) {
    const int s = (int)(k % STAGES);
    const long c = blockIdx.x + k * gridDim.x;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar[s])), "r"(CHUNK) : "memory");
    bulk_g2s(smem + (size_t)s * CHUNK, src + c * CHUNK, CHUNK, &bar[s]);
  };
  for (long k = 0; k < STAGES - 1 && k < mine; ++k) issue(k);
  for (long k = 0; k < mine; ++k) {
    const int s = (int)(k % STAGES);
    if (k + STAGES - 1 < mine) {
      // the buffer about to be refilled was stored STAGES-1 iterations ago: wait until its bulk store has read it
      asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(0) : "memory");
      issue(k + STAGES - 1);
    }
    mbar_wait(&bar[s], (unsigned)((k / STAGES) & 1));
    const long c = blockIdx.x + k * gridDim.x;
    bulk_s2g(dst + c * CHUNK, smem + (size_t)s * CHUNK, CHUNK);
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  }
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

template <typename F>
void bench(const char* name, F launch, double bytes, int dev0, int dev1) {
  for (int i = 0; i < 3; ++i) launch();
  CK(cudaSetDevice(dev0)); CK(cudaDeviceSynchronize());
  CK(cudaSetDevice(dev1)); CK(cudaDeviceSynchronize());
  CK(cudaSetDevice(dev0));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int reps = 20;
  cudaEventRecord(e0);
  for (int i = 0; i < reps; ++i) launch();
  cudaEventRecord(e1);
  CK(cudaEventSynchronize(e1));
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const float us = ms * 1e3f / reps;
  printf("%-58s %8.1f us  %7.1f GB/s per direction\n", name, us, bytes / us / 1e3);
}

int main() {
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  if (ndev < 2) { printf("needs 2 GPUs\n"); return 0; }
  const long N = 25583616;             // floats: 102.3 MB
  const long n8 = N / 8;
  const double B = 4.0 * N;
  float *a0, *b0, *a1, *b1;
  CK(cudaSetDevice(0)); CK(cudaDeviceEnablePeerAccess(1, 0)); CK(cudaMalloc(&a0, N * 4)); CK(cudaMalloc(&b0, N * 4));
  CK(cudaMemset(a0, 0, N * 4)); CK(cudaMemset(b0, 0, N * 4));
  CK(cudaSetDevice(1)); CK(cudaDeviceEnablePeerAccess(0, 0)); CK(cudaMalloc(&a1, N * 4)); CK(cudaMalloc(&b1, N * 4));
  CK(cudaMemset(a1, 0, N * 4)); CK(cudaMemset(b1, 0, N * 4));
  CK(cudaSetDevice(0));
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaStream_t s1, s2;   // non-blocking: the legacy default stream would serialise them
  CK(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
  cudaEvent_t fork, j1, j2;
  cudaEventCreateWithFlags(&fork, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&j1, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&j2, cudaEventDisableTiming);
  auto two = [&](auto fa, auto fb) {   // run fa on s1 and fb on s2 concurrently, fork/join on the default stream
    cudaEventRecord(fork, 0);
    cudaStreamWaitEvent(s1, fork, 0);
    cudaStreamWaitEvent(s2, fork, 0);
    fa(s1);
    fb(s2);
    cudaEventRecord(j1, s1);
    cudaEventRecord(j2, s2);
    cudaStreamWaitEvent(0, j1, 0);
    cudaStreamWaitEvent(0, j2, 0);
  };

  bench("cudaMemcpyPeerAsync 1 -> 0 (copy engine)", [&] { CK(cudaMemcpyPeerAsync(b0, 0, a1, 1, N * 4, 0)); }, B, 0, 1);
  bench("(a) pull LDG.256, grid 2/SM x512", [&] { pull_ldg<<<sms * 2, 512>>>(a1, b0, n8); }, B, 0, 1);
  bench("(a) pull LDG.256, grid 4/SM x512", [&] { pull_ldg<<<sms * 4, 512>>>(a1, b0, n8); }, B, 0, 1);
  bench("(c) push STG.256, grid 2/SM x512", [&] { push_stg<<<sms * 2, 512>>>(a0, b1, n8); }, B, 0, 1);
  bench("(c) push STG.256, grid 1/SM x512", [&] { push_stg<<<sms, 512>>>(a0, b1, n8); }, B, 0, 1);
  bench("pull + push fused per thread (LDG/STG), grid 2/SM", [&] { both_ldst<<<sms * 2, 512>>>(a1, b0, a0, b1, n8); }, B, 0, 1);
  bench("pull || push on two streams (LDG / STG)", [&] {
    two([&](cudaStream_t s) { pull_ldg<<<sms, 512, 0, s>>>(a1, b0, n8); },
        [&](cudaStream_t s) { push_stg<<<sms, 512, 0, s>>>(a0, b1, n8); });
  }, B, 0, 1);
  bench("split kernel: even blocks pull, odd blocks push", [&] { split_ldst<<<sms * 2, 512>>>(a1, b0, a0, b1, n8); }, B, 0, 1);
  bench("split kernel, 4 blocks/SM", [&] { split_ldst<<<sms * 4, 512>>>(a1, b0, a0, b1, n8); }, B, 0, 1);
  {
    constexpr int CH = 16384, ST = 4;
    auto k = copy_tma<CH, ST>;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, CH * ST));
    const long nch = N * 4 / CH;       // 6246 chunks
    bench("(b) pull TMA bulk 16 KB x4 stages, 1 CTA/SM", [&] { k<<<sms, 32, CH * ST>>>((const char*)a1, (char*)b0, nch); }, B, 0, 1);
    bench("(b) pull TMA bulk 16 KB x4 stages, 2 CTA/SM", [&] { k<<<sms * 2, 32, CH * ST>>>((const char*)a1, (char*)b0, nch); }, B, 0, 1);
    bench("(d) push TMA bulk 16 KB x4 stages, 1 CTA/SM", [&] { k<<<sms, 32, CH * ST>>>((const char*)a0, (char*)b1, nch); }, B, 0, 1);
    bench("(d) push TMA bulk 16 KB x4 stages, 2 CTA/SM", [&] { k<<<sms * 2, 32, CH * ST>>>((const char*)a0, (char*)b1, nch); }, B, 0, 1);
    bench("pull || push on two streams (TMA / TMA)", [&] {
      two([&](cudaStream_t s) { k<<<sms, 32, CH * ST, s>>>((const char*)a1, (char*)b0, nch); },
          [&](cudaStream_t s) { k<<<sms, 32, CH * ST, s>>>((const char*)a0, (char*)b1, nch); });
    }, B, 0, 1);
    bench("pull TMA || push STG on two streams", [&] {
      two([&](cudaStream_t s) { k<<<sms, 32, CH * ST, s>>>((const char*)a1, (char*)b0, nch); },
          [&](cudaStream_t s) { push_stg<<<sms, 512, 0, s>>>(a0, b1, n8); });
    }, B, 0, 1);
    bench("pull LDG || copy-engine push (cudaMemcpyPeerAsync)", [&] {
      two([&](cudaStream_t s) { pull_ldg<<<sms * 2, 512, 0, s>>>(a1, b0, n8); },
          [&](cudaStream_t s) { CK(cudaMemcpyPeerAsync(b1, 1, a0, 0, N * 4, s)); });
    }, B, 0, 1);
  }
  {
    constexpr int CH = 4096, ST = 8;
    auto k = copy_tma<CH, ST>;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, CH * ST));
    const long nch = N * 4 / CH;
    bench("(b) pull TMA bulk 4 KB x8 stages, 4 CTA/SM", [&] { k<<<sms * 4, 32, CH * ST>>>((const char*)a1, (char*)b0, nch); }, B, 0, 1);
    bench("(d) push TMA bulk 4 KB x8 stages, 4 CTA/SM", [&] { k<<<sms * 4, 32, CH * ST>>>((const char*)a0, (char*)b1, nch); }, B, 0, 1);
  }
  CK(cudaDeviceSynchronize());
  return 0;
}
