// apply_variants.cu -- developer microbenchmark (not part of the library): access-pattern
// variants of the N=1 fused Adam apply stream (4 reads + 3 writes of 102 MB each), timed
// back to back with CUDA events, to find what limits the kernel in steady state.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o apply_variants apply_variants.cu
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                     \
  do {                                                                            \
    cudaError_t e = (x);                                                          \
    if (e != cudaSuccess) {                                                       \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
      exit(1);                                                                    \
    }                                                                             \
  } while (0)

struct P {
  float alpha, omb1, omb2, eps;
};

__device__ __forceinline__ void adam1(float g, float& w, float& m, float& v, const P& o) {
  float d1 = __fsub_rn(g, m);
  m = __fadd_rn(m, __fmul_rn(d1, o.omb1));
  float d2 = __fsub_rn(__fmul_rn(g, g), v);
  v = __fadd_rn(v, __fmul_rn(d2, o.omb2));
  float num = __fmul_rn(o.alpha, m);
  float den = __fadd_rn(__fsqrt_rn(v), o.eps);
  w = __fsub_rn(w, __fdiv_rn(num, den));
}

template <int HINT>
__device__ __forceinline__ float4 ld4(const float4* p) {
  float4 v;
  if (HINT == 0)
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  else if (HINT == 1)
    asm volatile("ld.global.L1::no_allocate.L2::256B.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  else
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void st4(float4* p, float4 v) {
  asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// MATH 1: Adam; 0: trivial add (pure stream).  U vectors in flight per thread.
template <int THREADS, int U, int HINT, int MATH>
__global__ void __launch_bounds__(THREADS) apply_k(float* w, float* m, float* v, const float* g, long nvec, P o) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long base = (long)blockIdx.x * blockDim.x + threadIdx.x; base < nvec; base += stride * U) {
    float4 gg[U], ww[U], mm[U], vv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      long q = base + u * stride;
      if (q < nvec) {
        gg[u] = ld4<HINT>((const float4*)g + q);
        ww[u] = ld4<HINT>((const float4*)w + q);
        mm[u] = ld4<HINT>((const float4*)m + q);
        vv[u] = ld4<HINT>((const float4*)v + q);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      long q = base + u * stride;
      if (q < nvec) {
        if (MATH) {
          adam1(gg[u].x, ww[u].x, mm[u].x, vv[u].x, o);
          adam1(gg[u].y, ww[u].y, mm[u].y, vv[u].y, o);
          adam1(gg[u].z, ww[u].z, mm[u].z, vv[u].z, o);
          adam1(gg[u].w, ww[u].w, mm[u].w, vv[u].w, o);
        } else {
          mm[u].x += gg[u].x; mm[u].y += gg[u].y; mm[u].z += gg[u].z; mm[u].w += gg[u].w;
          vv[u].x += gg[u].x; vv[u].y += gg[u].y; vv[u].z += gg[u].z; vv[u].w += gg[u].w;
          ww[u].x += gg[u].x; ww[u].y += gg[u].y; ww[u].z += gg[u].z; ww[u].w += gg[u].w;
        }
        st4((float4*)m + q, mm[u]);
        st4((float4*)v + q, vv[u]);
        st4((float4*)w + q, ww[u]);
      }
    }
  }
}

// contiguous-chunk variant: a block owns CHUNK consecutive vectors of each stream at a time
template <int THREADS, int U, int MATH>
__global__ void __launch_bounds__(THREADS) apply_chunk_k(float* w, float* m, float* v, const float* g, long nvec, P o) {
  const long chunk = (long)THREADS * U;
  for (long c0 = (long)blockIdx.x * chunk; c0 < nvec; c0 += (long)gridDim.x * chunk) {
    float4 gg[U], ww[U], mm[U], vv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      long q = c0 + u * THREADS + threadIdx.x;
      if (q < nvec) {
        gg[u] = ld4<0>((const float4*)g + q);
        ww[u] = ld4<0>((const float4*)w + q);
        mm[u] = ld4<0>((const float4*)m + q);
        vv[u] = ld4<0>((const float4*)v + q);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      long q = c0 + u * THREADS + threadIdx.x;
      if (q < nvec) {
        if (MATH) {
          adam1(gg[u].x, ww[u].x, mm[u].x, vv[u].x, o);
          adam1(gg[u].y, ww[u].y, mm[u].y, vv[u].y, o);
          adam1(gg[u].z, ww[u].z, mm[u].z, vv[u].z, o);
          adam1(gg[u].w, ww[u].w, mm[u].w, vv[u].w, o);
        }
        st4((float4*)m + q, mm[u]);
        st4((float4*)v + q, vv[u]);
        st4((float4*)w + q, ww[u]);
      }
    }
  }
}

// 256-bit accesses (LDG.256 / STG.256)
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
template <int THREADS, int MATH>
__global__ void __launch_bounds__(THREADS) apply_256_k(float* w, float* m, float* v, const float* g, long nvec8, P o) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long q = (long)blockIdx.x * blockDim.x + threadIdx.x; q < nvec8; q += stride) {
    f8 gg = ld8(g + q * 8), ww = ld8(w + q * 8), mm = ld8(m + q * 8), vv = ld8(v + q * 8);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MATH) adam1(gg.a[i], ww.a[i], mm.a[i], vv.a[i], o);
      else { mm.a[i] += gg.a[i]; vv.a[i] += gg.a[i]; ww.a[i] += gg.a[i]; }
    }
    st8(m + q * 8, mm);
    st8(v + q * 8, vv);
    st8(w + q * 8, ww);
  }
}

// ---- TMA-staged variant: cp.async.bulk (1-D TMA) global->shared with an mbarrier, compute in
// shared memory, cp.async.bulk shared->global.  One tile of each stream per CTA; several CTAs
// per SM overlap load / compute / store.  Built to answer "would TMA/shared-memory staging
// beat plain 256-bit loads for this 7-stream pass?" with a measurement.
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}
template <int TILE, int MATH>
__global__ void __launch_bounds__(256) apply_tma_k(float* w, float* m, float* v, const float* g, long n, P o) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* sg = reinterpret_cast<float*>(smem_raw);
  float* sw = sg + TILE;
  float* sm = sw + TILE;
  float* sv = sm + TILE;
  __shared__ __align__(8) unsigned long long bar;
  const long e0 = (long)blockIdx.x * TILE;
  if (e0 >= n) return;
  const unsigned bytes = TILE * 4;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(4 * bytes) : "memory");
    bulk_g2s(sg, g + e0, bytes, &bar);
    bulk_g2s(sw, w + e0, bytes, &bar);
    bulk_g2s(sm, m + e0, bytes, &bar);
    bulk_g2s(sv, v + e0, bytes, &bar);
  }
  {
    unsigned done = 0;
    while (!done)
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                   : "=r"(done)
                   : "r"(smem_u32(&bar))
                   : "memory");
  }
#pragma unroll
  for (int j = 0; j < TILE / (4 * 256); ++j) {
    const int q = j * 256 + threadIdx.x;
    float4 gg = reinterpret_cast<float4*>(sg)[q], ww = reinterpret_cast<float4*>(sw)[q];
    float4 mm = reinterpret_cast<float4*>(sm)[q], vv = reinterpret_cast<float4*>(sv)[q];
    if (MATH) {
      adam1(gg.x, ww.x, mm.x, vv.x, o); adam1(gg.y, ww.y, mm.y, vv.y, o);
      adam1(gg.z, ww.z, mm.z, vv.z, o); adam1(gg.w, ww.w, mm.w, vv.w, o);
    } else {
      mm.x += gg.x; vv.x += gg.x; ww.x += gg.x;
    }
    reinterpret_cast<float4*>(sw)[q] = ww;
    reinterpret_cast<float4*>(sm)[q] = mm;
    reinterpret_cast<float4*>(sv)[q] = vv;
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    bulk_s2g(m + e0, sm, bytes);
    bulk_s2g(v + e0, sv, bytes);
    bulk_s2g(w + e0, sw, bytes);
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
}

__global__ void copy_k(const float4* a, float4* b, long nvec) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long q = (long)blockIdx.x * blockDim.x + threadIdx.x; q < nvec; q += stride) st4(b + q, ld4<0>(a + q));
}

__global__ void copy256_k(const float* a, float* b, long nvec8) {   // non-persistent, one 256-bit vector per thread
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q < nvec8) st8(b + q * 8, ld8(a + q * 8));
}
__global__ void copy256x4_k(const float* a, float* b, long nvec8) {  // one block = 8192 elements, 4 vectors per thread
  long base = (long)blockIdx.x * 1024 + threadIdx.x;
  f8 v[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) if (base + j * 256 < nvec8) v[j] = ld8(a + (base + j * 256) * 8);
#pragma unroll
  for (int j = 0; j < 4; ++j) if (base + j * 256 < nvec8) st8(b + (base + j * 256) * 8, v[j]);
}

template <typename F>
float bench(const char* name, F launch, double bytes) {
  for (int i = 0; i < 5; ++i) launch();
  CK(cudaDeviceSynchronize());
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int reps = 50;
  cudaEventRecord(e0);
  for (int i = 0; i < reps; ++i) launch();
  cudaEventRecord(e1);
  CK(cudaEventSynchronize(e1));
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  float us = ms * 1e3f / reps;
  printf("%-44s %8.1f us  %8.1f GB/s\n", name, us, bytes / us / 1e3);
  return us;
}

int main() {
  const long N = 25583616;  // padded ResNet-50 element count
  const long nvec = N / 4;
  float *w, *m, *v, *g, *c;
  CK(cudaMalloc(&w, N * 4)); CK(cudaMalloc(&m, N * 4)); CK(cudaMalloc(&v, N * 4)); CK(cudaMalloc(&g, N * 4));
  CK(cudaMalloc(&c, (size_t)1 << 30));
  CK(cudaMemset(w, 0, N * 4)); CK(cudaMemset(m, 0, N * 4)); CK(cudaMemset(v, 0, N * 4)); CK(cudaMemset(g, 0, N * 4));
  P o{3.16e-4f, 0.1f, 0.001f, 1e-7f};
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const double B = 28.0 * 25583592;
  printf("SMs %d\n", sms);
  bench("copy 512MB->512MB (1R:1W)", [&] { copy_k<<<sms * 8, 256>>>((const float4*)c, (float4*)(c + (1 << 27)), 1 << 25); }, 2.0 * (1 << 29));
  {
    const long n8 = N / 8;   // the pack-sized copy: 102 MB -> 102 MB
    const double CB = 8.0 * 25583592;
    bench("copy 102MB 256-bit non-persistent 1 vec/thr", [&] { copy256_k<<<(int)((n8 + 255) / 256), 256>>>(g, w, n8); }, CB);
    bench("copy 102MB 256-bit 4 vec/thr (pack shape)", [&] { copy256x4_k<<<(int)((n8 + 1023) / 1024), 256>>>(g, w, n8); }, CB);
    bench("copy 102MB 128-bit persistent 8/SM", [&] { copy_k<<<sms * 8, 256>>>((const float4*)g, (float4*)w, N / 4); }, CB);
    bench("copy 512MB 256-bit non-persistent 1 vec/thr", [&] { copy256_k<<<(1 << 24) / 256, 256>>>(c, c + (1 << 27), 1 << 24); }, 2.0 * (1 << 29));
  }
#define RUN(NAME, KERNEL, GRID, THREADS) bench(NAME, [&] { KERNEL<<<GRID, THREADS>>>(w, m, v, g, nvec, o); }, B)
  RUN("adam  512thr U2 grid 2/SM (library)", (apply_k<512, 2, 0, 1>), sms * 2, 512);
  RUN("adam  512thr U2 +L2::256B", (apply_k<512, 2, 1, 1>), sms * 2, 512);
  RUN("adam  512thr U2 ld.nc", (apply_k<512, 2, 2, 1>), sms * 2, 512);
  RUN("adam  256thr U2 grid 4/SM", (apply_k<256, 2, 0, 1>), sms * 4, 256);
  RUN("adam  256thr U4 grid 2/SM", (apply_k<256, 4, 0, 1>), sms * 2, 256);
  RUN("adam  512thr U4 grid 1/SM", (apply_k<512, 4, 0, 1>), sms * 1, 512);
  RUN("adam  1024thr U1 grid 1/SM", (apply_k<1024, 1, 0, 1>), sms * 1, 1024);
  RUN("adam  512thr U1 grid 2/SM", (apply_k<512, 1, 0, 1>), sms * 2, 512);
  RUN("adam  512thr U2 non-persistent grid", (apply_k<512, 2, 0, 1>), (int)((nvec + 1023) / 1024), 512);
  RUN("adam  256thr U2 non-persistent grid", (apply_k<256, 2, 0, 1>), (int)((nvec + 511) / 512), 256);
  RUN("adam  chunked 512thr U2 grid 2/SM", (apply_chunk_k<512, 2, 1>), sms * 2, 512);
  RUN("adam  chunked 256thr U4 grid 4/SM", (apply_chunk_k<256, 4, 1>), sms * 4, 256);
  RUN("adam  chunked 512thr U2 one chunk/block", (apply_chunk_k<512, 2, 1>), (int)((nvec + 1023) / 1024), 512);
  RUN("strm  512thr U2 grid 2/SM (no math)", (apply_k<512, 2, 0, 0>), sms * 2, 512);
  RUN("strm  512thr U2 +L2::256B (no math)", (apply_k<512, 2, 1, 0>), sms * 2, 512);
  RUN("strm  256thr U4 grid 4/SM (no math)", (apply_k<256, 4, 0, 0>), sms * 4, 256);
  RUN("strm  512thr U2 non-persistent (no math)", (apply_k<512, 2, 0, 0>), (int)((nvec + 1023) / 1024), 512);
  bench("adam  256-bit 256thr grid 4/SM", [&] { apply_256_k<256, 1><<<sms * 4, 256>>>(w, m, v, g, N / 8, o); }, B);
  bench("adam  256-bit 512thr grid 2/SM", [&] { apply_256_k<512, 1><<<sms * 2, 512>>>(w, m, v, g, N / 8, o); }, B);
  bench("adam  256-bit 256thr non-persistent", [&] { apply_256_k<256, 1><<<(int)((N / 8 + 255) / 256), 256>>>(w, m, v, g, N / 8, o); }, B);
  {
    auto k2 = apply_tma_k<2048, 1>;
    auto k4 = apply_tma_k<4096, 1>;
    auto k8 = apply_tma_k<8192, 1>;
    auto k4s = apply_tma_k<4096, 0>;
    CK(cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 2048 * 4));
    CK(cudaFuncSetAttribute(k4, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 4096 * 4));
    CK(cudaFuncSetAttribute(k8, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 8192 * 4));
    CK(cudaFuncSetAttribute(k4s, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 4096 * 4));
    bench("adam  TMA bulk tile 2048 (32 KB smem/CTA)", [&] { k2<<<(int)(N / 2048), 256, 4 * 2048 * 4>>>(w, m, v, g, N, o); }, B);
    bench("adam  TMA bulk tile 4096 (64 KB smem/CTA)", [&] { k4<<<(int)(N / 4096), 256, 4 * 4096 * 4>>>(w, m, v, g, N, o); }, B);
    bench("adam  TMA bulk tile 8192 (128 KB smem/CTA)", [&] { k8<<<(int)(N / 8192), 256, 4 * 8192 * 4>>>(w, m, v, g, N, o); }, B);
    bench("strm  TMA bulk tile 4096 (no math)", [&] { k4s<<<(int)(N / 4096), 256, 4 * 4096 * 4>>>(w, m, v, g, N, o); }, B);
    CK(cudaDeviceSynchronize());
  }
  bench("strm  256-bit 256thr grid 4/SM (no math)", [&] { apply_256_k<256, 0><<<sms * 4, 256>>>(w, m, v, g, N / 8, o); }, B);
  return 0;
}
