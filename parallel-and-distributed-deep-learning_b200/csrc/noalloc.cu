// noalloc.cu -- b2dp_adam_apply_noalloc: the fused Adam apply + gradient statistics for
// callers that are not allowed to allocate device memory.
//
// Who that is: code that runs where cudaMalloc is off limits -- inside a framework op
// whose device memory must all come from the framework's own allocator (a TensorFlow or
// Horovod custom-op Compute(), a torch autograd hook under the caching allocator's
// accounting), inside a latency-critical loop where an allocation would serialise the
// device, or in a process whose memory budget is fixed up front.  Those callers already
// own w, m, v and g; what they cannot supply is the few KB of scratch the statistics
// reduction needs.  Here that scratch is a file-scope __device__ array that every GPU
// loading the library carries until the process exits (kScratchBytes, 128.25 KB), so the
// entry point never calls cudaMalloc/cudaFree or any allocating routine.
//
// The price of a fixed array is a fixed maximum: one partial sum per 8192-element chunk
// and 8192 chunk slots give B2DP_NOALLOC_MAX_ELEMS = 2^26 elements per call (2.6x the
// 25,583,592 of ResNet-50); larger calls are refused with B2DP_ERR_TOO_LARGE.  All calls
// on a device share the array: two calls that overlap on the GPU (different streams, no
// ordering between them) race on it and corrupt each other's statistics.
#include <algorithm>
#include <cstring>
#include <mutex>

#include "kernels.cuh"

using namespace b2dp;

namespace {

constexpr int kChunkElems = 8192;                      // one partial (sum g^2, non-finite count) per chunk
constexpr int kChunkVecs = kChunkElems / 4;
constexpr int kMaxChunks = 8192;                       // => 2^26 elements per call
constexpr size_t kTicketBytes = 256;                   // arrival counter, padded to the array's alignment
constexpr size_t kScratchBytes = kTicketBytes + size_t(kMaxChunks) * 2 * sizeof(double);
constexpr int kMaxDevices = 64;

static_assert(int64_t(kMaxChunks) * kChunkElems == B2DP_NOALLOC_MAX_ELEMS, "header and scratch size disagree");

// The scratch itself.  In host code this name is only a handle for cudaGetSymbolAddress;
// it is NOT a device address and must never be handed to a kernel as one.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

// Same arithmetic as step_kernel<1, float, ADAM> (adam_update: one rounding per operation),
// but the statistics are reduced per fixed 8192-element chunk and folded in chunk order,
// so they do not depend on the grid the kernel happens to be launched with.
__global__ void __launch_bounds__(kStepThreads)
adam_apply_chunked_kernel(float* __restrict__ w, float* __restrict__ m, float* __restrict__ v,
                          const float* __restrict__ g, int64_t numel, int64_t nchunks, OptScalars o,
                          double* __restrict__ partials, uint32_t* __restrict__ ticket,
                          double* __restrict__ stats_out) {
  __shared__ double sm[2][kStepThreads / 32];
  __shared__ bool last;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t nvec = numel >> 2;
  for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    float sumsq = 0.0f, bad = 0.0f;
    const int64_t v0 = c * kChunkVecs;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      float4 gg[2], ww[2], mm[2], vv[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int64_t q = v0 + (half * 2 + u) * kStepThreads + threadIdx.x;
        if (q < nvec) {
          gg[u] = ld_stream_f4(reinterpret_cast<const float4*>(g) + q);
          ww[u] = ld_stream_f4(reinterpret_cast<const float4*>(w) + q);
          mm[u] = ld_stream_f4(reinterpret_cast<const float4*>(m) + q);
          vv[u] = ld_stream_f4(reinterpret_cast<const float4*>(v) + q);
        }
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int64_t q = v0 + (half * 2 + u) * kStepThreads + threadIdx.x;
        if (q < nvec) {
          float4 a = gg[u];
          a.x = __fmul_rn(a.x, o.postscale);
          a.y = __fmul_rn(a.y, o.postscale);
          a.z = __fmul_rn(a.z, o.postscale);
          a.w = __fmul_rn(a.w, o.postscale);
          sumsq = fmaf(a.x, a.x, fmaf(a.y, a.y, fmaf(a.z, a.z, fmaf(a.w, a.w, sumsq))));
          bad += (float)(!isfinite(a.x)) + (float)(!isfinite(a.y)) + (float)(!isfinite(a.z)) +
                 (float)(!isfinite(a.w));
          adam_update(a.x, ww[u].x, mm[u].x, vv[u].x, o);
          adam_update(a.y, ww[u].y, mm[u].y, vv[u].y, o);
          adam_update(a.z, ww[u].z, mm[u].z, vv[u].z, o);
          adam_update(a.w, ww[u].w, mm[u].w, vv[u].w, o);
          st_f4(reinterpret_cast<float4*>(m) + q, mm[u]);
          st_f4(reinterpret_cast<float4*>(v) + q, vv[u]);
          st_f4(reinterpret_cast<float4*>(w) + q, ww[u]);
        }
      }
    }
    // the last chunk also owns the numel % 4 scalar tail
    if (c == nchunks - 1) {
      const int64_t i = (nvec << 2) + threadIdx.x;
      if (i < numel) {
        float a = __fmul_rn(g[i], o.postscale);
        sumsq = fmaf(a, a, sumsq);
        bad += (float)(!isfinite(a));
        float wi = w[i], mi = m[i], vi = v[i];
        adam_update(a, wi, mi, vi, o);
        w[i] = wi;
        m[i] = mi;
        v[i] = vi;
      }
    }
    double s = (double)sumsq, b = (double)bad;
#pragma unroll
    for (int k = 16; k > 0; k >>= 1) {
      s += __shfl_xor_sync(0xffffffffu, s, k);
      b += __shfl_xor_sync(0xffffffffu, b, k);
    }
    if (lane == 0) {
      sm[0][warp] = s;
      sm[1][warp] = b;
    }
    __syncthreads();
    if (warp == 0) {
      s = lane < kStepThreads / 32 ? sm[0][lane] : 0.0;
      b = lane < kStepThreads / 32 ? sm[1][lane] : 0.0;
#pragma unroll
      for (int k = 16; k > 0; k >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, k);
        b += __shfl_xor_sync(0xffffffffu, b, k);
      }
      if (lane == 0) {
        partials[2 * c] = s;
        partials[2 * c + 1] = b;
      }
    }
    __syncthreads();
  }
  // last block to arrive folds the chunk partials in chunk order
  if (threadIdx.x == 0) {
    __threadfence();
    last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (last && warp == 0) {
    __threadfence();
    double s = 0.0, b = 0.0;
    for (int64_t i = lane; i < nchunks; i += 32) {
      s += __ldcg(partials + 2 * i);
      b += __ldcg(partials + 2 * i + 1);
    }
#pragma unroll
    for (int k = 16; k > 0; k >>= 1) {
      s += __shfl_xor_sync(0xffffffffu, s, k);
      b += __shfl_xor_sync(0xffffffffu, b, k);
    }
    if (lane == 0) {
      stats_out[0] = s;
      stats_out[1] = b;
    }
  }
}

struct DeviceSetup {
  bool done = false;
  unsigned char* scratch = nullptr;  // device address of g_noalloc_scratch on this device
  int max_grid = 0;
};
DeviceSetup g_setup[kMaxDevices];
std::mutex g_setup_mu;

// Once per device, with that device current: resolve the device address of the scratch
// array (cudaGetSymbolAddress) and the resident grid size.  Allocates nothing.
int setup_current_device(int device, DeviceSetup** out) {
  std::lock_guard<std::mutex> lock(g_setup_mu);
  DeviceSetup& s = g_setup[device];
  if (!s.done) {
    int major = 0;
    cudaError_t e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
    if (e != cudaSuccess) return fail(B2DP_ERR_CUDA, "adam_apply_noalloc: cudaDeviceGetAttribute: %s", cudaGetErrorString(e));
    if (major != 10)
      return fail(B2DP_ERR_ARCH, "adam_apply_noalloc: device %d is sm_%dx; this library is built for sm_100a only", device, major);
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr) {
      (void)cudaGetLastError();
      return fail(B2DP_ERR_CUDA, "adam_apply_noalloc: cudaGetSymbolAddress(scratch) on device %d: %s", device,
                  cudaGetErrorString(e));
    }
    int sms = 0, occ = 0;
    e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (e == cudaSuccess)
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, adam_apply_chunked_kernel, kStepThreads, 0);
    if (e != cudaSuccess || sms < 1 || occ < 1) {
      (void)cudaGetLastError();
      return fail(B2DP_ERR_CUDA, "adam_apply_noalloc: occupancy query on device %d: %s", device, cudaGetErrorString(e));
    }
    s.scratch = static_cast<unsigned char*>(ptr);
    s.max_grid = sms * occ;
    s.done = true;
  }
  *out = &s;
  return B2DP_OK;
}

}  // namespace

extern "C" int b2dp_adam_apply_noalloc(float* w, float* m, float* v, const float* g, int64_t numel,
                                       const b2dp_hyper* hyper, double* stats_out, void* stream) {
  if (numel < 0) return fail(B2DP_ERR_INVALID, "adam_apply_noalloc: negative numel %lld", (long long)numel);
  if (numel > B2DP_NOALLOC_MAX_ELEMS)
    return fail(B2DP_ERR_TOO_LARGE,
                "adam_apply_noalloc: %lld elements exceed the fixed scratch (max %lld per call; split the call "
                "or use a plan with b2dp_step)",
                (long long)numel, (long long)B2DP_NOALLOC_MAX_ELEMS);
  if (!hyper || !stats_out) return fail(B2DP_ERR_INVALID, "adam_apply_noalloc: hyper and stats_out are required");
  if (hyper->step < 1) return fail(B2DP_ERR_INVALID, "adam_apply_noalloc: Adam step must be >= 1");
  if (numel > 0 && (!w || !m || !v || !g)) return fail(B2DP_ERR_INVALID, "adam_apply_noalloc: null tensor pointer");
  if ((reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v) |
       reinterpret_cast<uintptr_t>(g)) & 15)
    return fail(B2DP_ERR_INVALID, "adam_apply_noalloc: w, m, v and g must be 16-byte aligned");
  if (reinterpret_cast<uintptr_t>(stats_out) & 7)
    return fail(B2DP_ERR_INVALID, "adam_apply_noalloc: stats_out must be 8-byte aligned");

  int device = -1;
  cudaError_t e = cudaGetDevice(&device);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return fail(B2DP_ERR_NO_DEVICE, "adam_apply_noalloc: no CUDA device: %s (there is no CPU fallback)", cudaGetErrorString(e));
  }
  if (device < 0 || device >= kMaxDevices) return fail(B2DP_ERR_INVALID, "adam_apply_noalloc: device index %d unsupported", device);
  DeviceSetup* s = nullptr;
  if (int rc = setup_current_device(device, &s)) return rc;

  cudaStream_t st = (cudaStream_t)stream;
  if (numel == 0) {
    e = cudaMemsetAsync(stats_out, 0, 2 * sizeof(double), st);
    if (e != cudaSuccess) return fail(B2DP_ERR_CUDA, "adam_apply_noalloc: cudaMemsetAsync: %s", cudaGetErrorString(e));
    return B2DP_OK;
  }
  OptScalars o;
  memset(&o, 0, sizeof o);
  o.alpha = b2dp_adam_alpha(hyper->lr, hyper->beta1, hyper->beta2, hyper->step);
  o.omb1 = 1.0f - hyper->beta1;
  o.omb2 = 1.0f - hyper->beta2;
  o.eps = hyper->eps;
  o.postscale = hyper->postscale;

  uint32_t* ticket = reinterpret_cast<uint32_t*>(s->scratch);
  double* partials = reinterpret_cast<double*>(s->scratch + kTicketBytes);
  // The array keeps whatever the previous call (or an aborted one) left: the arrival
  // counter is the one word that must start at zero, so clear it in stream order.
  e = cudaMemsetAsync(ticket, 0, sizeof(uint32_t), st);
  if (e != cudaSuccess) return fail(B2DP_ERR_CUDA, "adam_apply_noalloc: cudaMemsetAsync(ticket): %s", cudaGetErrorString(e));
  const int64_t nchunks = (numel + kChunkElems - 1) / kChunkElems;
  const int grid = (int)std::min<int64_t>(nchunks, s->max_grid);
  adam_apply_chunked_kernel<<<grid, kStepThreads, 0, st>>>(w, m, v, g, numel, nchunks, o, partials, ticket, stats_out);
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail(B2DP_ERR_CUDA, "adam_apply_noalloc: launch: %s", cudaGetErrorString(e));
  return B2DP_OK;
}
