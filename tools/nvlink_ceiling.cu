// nvlink_ceiling.cu -- developer microbenchmark (not part of the library): what NVLink rate does the
// step kernel's traffic pattern get when ALL n GPUs of the box issue it at once?
//
// One process, n GPUs with peer access, S = 102.33 MB per GPU (the ResNet-50 gradient bucket), every
// GPU running the same kernel concurrently, exactly as the collective step does:
//   pull   : GPU r loads slice r (S/n) from each of the n-1 peers with LDG.256, adds, stores locally
//            (the reduce-scatter half)                      NVLink bytes per GPU per direction (n-1)/n*S
//   push   : GPU r stores its slice r into each of the n-1 peers with STG.256 (the all-gather half)
//                                                           (n-1)/n*S per direction
//   fused  : both in one kernel, per thread (what step256_kernel does)      2(n-1)/n*S per direction
//   pull2  : pull twice (reduce-scatter by pull, then all-gather by pull: every rank reads the n-1
//            foreign slices) -- tells whether an all-gather by PULL would beat the push
// Times are CUDA events per device around `reps` back-to-back launches, max over devices.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o nvlink_ceiling nvlink_ceiling.cu
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                        \
  do {                                                                               \
    cudaError_t e = (x);                                                             \
    if (e != cudaSuccess) {                                                          \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
      exit(1);                                                                       \
    }                                                                                \
  } while (0)

constexpr int kMaxN = 8;
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

struct Args {
  const float* src[kMaxN];  // slice `rank` inside every GPU's source buffer
  float* dst[kMaxN];        // slice `rank` inside every GPU's destination buffer
  long nvec8;
  int rank;
};

template <int N, bool PULL, bool PUSH>
__global__ void __launch_bounds__(512) traffic(const Args a) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long q = (long)blockIdx.x * blockDim.x + threadIdx.x; q < a.nvec8; q += stride) {
    f8 acc = ld8(a.src[a.rank] + q * 8);
    if (PULL) {
      f8 g[N];
#pragma unroll
      for (int r = 0; r < N; ++r)
        if (r != a.rank) g[r] = ld8(a.src[r] + q * 8);
#pragma unroll
      for (int r = 0; r < N; ++r)
        if (r != a.rank)
#pragma unroll
          for (int i = 0; i < 8; ++i) acc.a[i] += g[r].a[i];
    }
    if (PUSH) {
#pragma unroll
      for (int r = 0; r < N; ++r) st8(a.dst[r] + q * 8, acc);
    } else {
      st8(a.dst[a.rank] + q * 8, acc);
    }
  }
}

// all-gather by pull: every rank reads the n-1 foreign slices from their owners into its own buffer
template <int N>
__global__ void __launch_bounds__(512) gather_pull(const Args a, long slice_floats) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long q = (long)blockIdx.x * blockDim.x + threadIdx.x; q < a.nvec8; q += stride) {
#pragma unroll
    for (int r = 0; r < N; ++r)
      if (r != a.rank) {
        // slice r lives at offset (r - rank) * slice from this rank's slice pointer
        const long off = (long)(r - a.rank) * slice_floats + q * 8;
        st8(a.dst[a.rank] + off, ld8(a.src[r] + off));
      }
  }
}

template <int N>
void run(int blocks_per_sm, int reps, long total) {
  const long slice = total / N;
  std::vector<float*> src(N), dst(N);
  std::vector<cudaEvent_t> e0(N), e1(N);
  int sms = 0;
  for (int d = 0; d < N; ++d) {
    CK(cudaSetDevice(d));
    CK(cudaMalloc(&src[d], total * 4));
    CK(cudaMalloc(&dst[d], total * 4));
    CK(cudaMemset(src[d], 0, total * 4));
    CK(cudaMemset(dst[d], 0, total * 4));
    CK(cudaEventCreate(&e0[d]));
    CK(cudaEventCreate(&e1[d]));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, d));
  }
  auto args_for = [&](int d) {
    Args a;
    for (int r = 0; r < kMaxN; ++r) { a.src[r] = nullptr; a.dst[r] = nullptr; }
    for (int r = 0; r < N; ++r) {
      a.src[r] = src[r] + (long)d * slice;
      a.dst[r] = dst[r] + (long)d * slice;
    }
    a.nvec8 = slice / 8;
    a.rank = d;
    return a;
  };
  auto time_it = [&](const char* name, auto launch, double bytes_per_dir) {
    for (int w = 0; w < 3; ++w)
      for (int d = 0; d < N; ++d) { CK(cudaSetDevice(d)); launch(d); }
    for (int d = 0; d < N; ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); }
    for (int d = 0; d < N; ++d) { CK(cudaSetDevice(d)); CK(cudaEventRecord(e0[d])); }
    for (int i = 0; i < reps; ++i)
      for (int d = 0; d < N; ++d) { CK(cudaSetDevice(d)); launch(d); }
    for (int d = 0; d < N; ++d) { CK(cudaSetDevice(d)); CK(cudaEventRecord(e1[d])); }
    float worst = 0;
    for (int d = 0; d < N; ++d) {
      CK(cudaSetDevice(d));
      CK(cudaEventSynchronize(e1[d]));
      float ms;
      CK(cudaEventElapsedTime(&ms, e0[d], e1[d]));
      worst = std::max(worst, ms);
    }
    const double us = worst * 1e3 / reps;
    printf("{\"n\": %d, \"variant\": \"%s\", \"blocks_per_sm\": %d, \"us\": %.1f, \"nvlink_bytes_per_gpu_per_direction\": %.0f, "
           "\"gbs_per_direction\": %.1f}\n", N, name, blocks_per_sm, us, bytes_per_dir, bytes_per_dir / us / 1e3);
    fflush(stdout);
  };
  const int grid = sms * blocks_per_sm;
  const double half = (double)(N - 1) / N * total * 4.0;
  time_it("pull", [&](int d) { traffic<N, true, false><<<grid, 512>>>(args_for(d)); }, half);
  time_it("push", [&](int d) { traffic<N, false, true><<<grid, 512>>>(args_for(d)); }, half);
  time_it("fused_pull_push", [&](int d) { traffic<N, true, true><<<grid, 512>>>(args_for(d)); }, 2 * half);
  time_it("pull_then_push_two_launches", [&](int d) {
    traffic<N, true, false><<<grid, 512>>>(args_for(d));
    traffic<N, false, true><<<grid, 512>>>(args_for(d));
  }, 2 * half);
  time_it("pull_then_gather_by_pull", [&](int d) {
    traffic<N, true, false><<<grid, 512>>>(args_for(d));
    gather_pull<N><<<grid, 512>>>(args_for(d), slice);
  }, 2 * half);
  time_it("gather_by_pull", [&](int d) { gather_pull<N><<<grid, 512>>>(args_for(d), slice); }, half);
  for (int d = 0; d < N; ++d) {
    CK(cudaSetDevice(d));
    CK(cudaFree(src[d]));
    CK(cudaFree(dst[d]));
  }
}

int main(int argc, char** argv) {
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  const int reps = argc > 1 ? atoi(argv[1]) : 20;
  for (int d = 0; d < ndev && d < kMaxN; ++d) {
    CK(cudaSetDevice(d));
    for (int p = 0; p < ndev && p < kMaxN; ++p)
      if (p != d) {
        cudaError_t e = cudaDeviceEnablePeerAccess(p, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
          printf("cudaDeviceEnablePeerAccess(%d -> %d): %s\n", d, p, cudaGetErrorString(e));
          return 1;
        }
        (void)cudaGetLastError();
      }
  }
  const long total = 25583616;  // floats; divisible by 8 * 8 * 8
  for (int bps : {1, 2}) {
    if (ndev >= 2) run<2>(bps, reps, total);
    if (ndev >= 4) run<4>(bps, reps, total);
    if (ndev >= 8) run<8>(bps, reps, total);
  }
  return 0;
}
