// api.cu -- contexts, plans and kernel launchers behind include/b200dp.h.
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>

#include <unistd.h>

#include <nvtx3/nvToolsExt.h>  // header-only; does nothing unless a profiler injects itself

#include "kernels.cuh"

using namespace b2dp;

namespace {

#define CUDA_TRY(call)                                                                   \
  do {                                                                                   \
    cudaError_t _e = (call);                                                             \
    if (_e != cudaSuccess) {                                                             \
      (void)cudaGetLastError();                                                          \
      return fail(B2DP_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(_e));               \
    }                                                                                    \
  } while (0)

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// NVTX range around every hot-path entry point, so an nsys/ncu timeline shows pack / step /
// all-reduce as named host ranges (the reference has no tracing of its own).
struct Range {
  explicit Range(const char* name) { nvtxRangePushA(name); }
  ~Range() { nvtxRangePop(); }
};

// Byte map of a rank's peer-mappable block.
struct BlockMap {
  size_t flags = 0;      // 2 phases * kMaxBlocks * kMaxRanks u32
  size_t ctrl = 0;       // status word, ticket, stats
  size_t partials = 0;   // 2 doubles per block
  size_t small = 0;      // small-message all-reduce slots: 2 halves x kMaxRanks x kSmallMax f32
  size_t grads = 0, weights = 0, reduced = 0, s0 = 0, s1 = 0, end = 0;
};

constexpr size_t kCtrlStatus = 0, kCtrlTicket = 64, kCtrlStats = 128;

BlockMap make_map(const b2dp_layout& l) {
  BlockMap m;
  const size_t wire = l.wire == B2DP_BF16 ? 2 : 4;
  size_t off = 0;
  m.flags = off;
  off += align_up(size_t(2) * kMaxBlocks * kMaxRanks * sizeof(uint32_t), 4096);
  m.ctrl = off;
  off += 4096;
  m.partials = off;
  off += align_up(size_t(2) * kApplyMaxBlocks * sizeof(double), 4096);
  m.small = off;
  off += align_up(size_t(2) * kMaxRanks * kSmallMax * sizeof(float), 4096);
  m.grads = off;
  off += align_up((size_t)l.total * wire, 4096);
  m.weights = off;
  off += align_up((size_t)l.total * 4, 4096);
  if (l.wire == B2DP_F32) {
    m.reduced = m.grads;  // the all-reduce runs in place
  } else {
    m.reduced = off;
    off += align_up((size_t)l.total * 4, 4096);
  }
  m.s0 = off;
  off += align_up((size_t)l.shard * 4, 4096);
  m.s1 = off;
  off += align_up((size_t)l.shard * 4, 4096);
  m.end = off;
  return m;
}

// A caller's K-pointer array as the kernels use it: the plan's tile table with those pointers
// resolved (TileAbs), rebuilt by one small kernel only when the array changes.  Everything is
// allocated when the plan is created; the hot path never allocates.
struct PtrTable {
  std::vector<const void*> host;   // the array the resolved table was built from
  TileAbs* abs = nullptr;          // one entry per tile of the layout
  bool valid = false;
  void** dev = nullptr;            // device copy of the array; only for K > kParamPtrs
};

struct RankState {
  int device = -1;
  int rank = -1;
  VmmBlock own;
  VmmBlock imported[kMaxRanks];
  char* base[kMaxRanks] = {};  // every rank's block as addressable from this rank's GPU
  bool attached[kMaxRanks] = {};
  Tile* d_tiles = nullptr;
  PtrTable grads, outs, weights;
  uint32_t epoch = 0;
  int num_sms = 0;
  // verdict of the cross-GPU flag waits in pinned, mapped host memory: a kernel that gives up
  // writes it, and the next collective call of this rank fails fast on it without a sync
  int32_t* host_status = nullptr;      // host address
  int32_t* host_status_dev = nullptr;  // the same word as the GPU addresses it
  // multi-bucket pack_step: bucket b+1 is packed on `side` while bucket b's collective step runs
  cudaStream_t side = nullptr;
  cudaEvent_t ev_start = nullptr;
  std::vector<cudaEvent_t> ev_packed;  // one per bucket
  // Path B: the plan's multicast object as seen by this rank
  unsigned long long mc_handle = 0, mc_va = 0;
  bool mc_owner = false, mc_joined = false, mc_bound = false;
  uint32_t small_calls = 0;          // parity selects the slot half
  float* small_out = nullptr;        // emulated ranks: the sum phase is deferred to the last rank's call
  int32_t small_count = 0;
  float small_post = 1.0f;
};

}  // namespace

struct b2dp_ctx {
  bool local = true;
  bool emulated = false;  // same device listed more than once
  int nranks = 1;
  int nlocal = 1;
  int devices[kMaxRanks] = {};
  int ranks[kMaxRanks] = {};
  double timeout_s = 10.0;
};

struct b2dp_plan {
  b2dp_ctx* ctx = nullptr;
  b2dp_layout layout;
  BlockMap map;
  std::vector<std::unique_ptr<RankState>> rs;
  size_t block_bytes = 0;  // allocated size of every rank's block (granularity-rounded)
  bool ctx_local = true;   // copy of ctx->local: destroy must not touch a context that died first
  int path = B2DP_PATH_PEER;
  // launch geometry knobs (b2dp_plan_set_tuning); defaults from the r01 sweeps on 8xB200
  int grid_cap = kMaxBlocks;  // peer-path blocks (also bounded by co-residency)
  int mc_grid_cap = 74;       // multicast path: fewer blocks contend less in the switch
  int mc_unroll = 1;
  int mc_bf16_grid_cap = 148; // multicast path, bf16 wire (8 elements per request)
  int skip_own = 0;           // pack_step, peer path: leave the rank's own slice unpacked and read it in place
                              // (opt-in: measured no faster)
  int peer_unroll = 0;        // 0 = automatic (1 for n == 8, else 2)
  int peer_vec = 8;           // elements per peer access: 8 (256-bit; f32 wire, n in {2,4,8}) or 4 (128-bit)
};

namespace {

int check_device(int device) {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    (void)cudaGetLastError();
    return fail(B2DP_ERR_NO_DEVICE, "no CUDA device: %s (libb200dp has no CPU fallback)",
                e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  }
  if (device < 0 || device >= count) return fail(B2DP_ERR_INVALID, "device %d outside 0..%d", device, count - 1);
  int major = 0;
  CUDA_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  if (major != 10) return fail(B2DP_ERR_ARCH, "device %d is sm_%dx; this library is built for sm_100a only", device, major);
  return B2DP_OK;
}

// Brings `t.abs` up to date with the caller's pointer array.  Unchanged array (the common case:
// frameworks keep gradient buffers at stable addresses): nothing is launched.  Changed: one
// resolve kernel; for up to kParamPtrs tensors the pointers ride in its launch parameters.
int sync_tiles(const b2dp_layout& l, Tile* d_tiles, PtrTable& t, const void* const* ptrs, cudaStream_t s) {
  const int K = l.K;
  const int64_t ntiles = (int64_t)l.tiles.size();
  if (K == 0 || ntiles == 0) return B2DP_OK;
  if (t.valid && (int)t.host.size() == K && memcmp(t.host.data(), ptrs, sizeof(void*) * K) == 0) return B2DP_OK;
  t.host.assign(ptrs, ptrs + K);
  const unsigned grid = (unsigned)((ntiles + 255) / 256);
  if (K <= kParamPtrs) {
    PtrPack pk;
    memcpy(pk.p, ptrs, sizeof(void*) * K);
    resolve_tiles_param_kernel<<<grid, 256, 0, s>>>(d_tiles, ntiles, pk, t.abs);
  } else {
    // pageable source: the runtime stages it before returning, so t.host may change later
    CUDA_TRY(cudaMemcpyAsync(t.dev, t.host.data(), sizeof(void*) * K, cudaMemcpyHostToDevice, s));
    resolve_tiles_kernel<<<grid, 256, 0, s>>>(d_tiles, ntiles, t.dev, t.abs);
  }
  CUDA_TRY(cudaGetLastError());
  t.valid = true;
  return B2DP_OK;
}

int get_rank(b2dp_plan* p, int dev, RankState** out) {
  if (!p) return fail(B2DP_ERR_INVALID, "plan is null");
  if (dev < 0 || dev >= (int)p->rs.size()) return fail(B2DP_ERR_INVALID, "dev %d outside 0..%d", dev, (int)p->rs.size() - 1);
  *out = p->rs[dev].get();
  return B2DP_OK;
}

// A flag wait of an earlier collective gave up (the kernel wrote the pinned host word): every
// later collective of this rank is refused until b2dp_plan_check has reported and cleared it,
// so training cannot run on with weights that were only partly updated.
int check_poisoned(const RankState& r) {
  if (r.host_status && *reinterpret_cast<volatile int32_t*>(r.host_status) != 0)
    return fail(B2DP_ERR_TIMEOUT,
                "rank %d: an earlier collective timed out waiting for a peer; the weights of this plan are not "
                "consistent across ranks any more (b2dp_plan_check reports and clears the condition)", r.rank);
  return B2DP_OK;
}

int bucket_range(const b2dp_plan* p, int bucket, int* b0, int* b1) {
  const int nb = (int)p->layout.buckets.size();
  if (bucket < 0) {
    *b0 = 0;
    *b1 = nb;
  } else if (bucket < nb) {
    *b0 = bucket;
    *b1 = bucket + 1;
  } else {
    return fail(B2DP_ERR_INVALID, "bucket %d outside 0..%d", bucket, nb - 1);
  }
  return B2DP_OK;
}

int pack_grid(const RankState&, int64_t ntiles) { return (int)ntiles; }  // one tile per block

template <int DIR>
int launch_pack(b2dp_plan* p, RankState& r, int bucket, PtrTable& table, const void* const* ptrs,
                size_t flat_off, int wire, float prescale, cudaStream_t s, int64_t skip_lo = 0, int64_t skip_hi = 0) {
  int b0 = 0, b1 = 0;
  if (int rc = bucket_range(p, bucket, &b0, &b1)) return rc;
  if (!ptrs && p->layout.K > 0) return fail(B2DP_ERR_INVALID, "tensor pointer array is null");
  for (int b = b0; b < b1; ++b) {
    const Bucket& bk = p->layout.buckets[b];
    if (bk.tile_count == 0) continue;
    for (int k = bk.first; k < bk.first + bk.count; ++k)
      if (p->layout.numel[k] > 0 && !ptrs[k]) return fail(B2DP_ERR_INVALID, "tensor %d pointer is null", k);
  }
  // buckets are adjacent in the tile table, so a range of buckets is one launch
  const int64_t t0 = p->layout.buckets.empty() ? 0 : p->layout.buckets[b0].tile_first;
  int64_t nt = 0;
  for (int b = b0; b < b1; ++b) nt += p->layout.buckets[b].tile_count;
  if (nt == 0) return B2DP_OK;
  if (int rc = sync_tiles(p->layout, r.d_tiles, table, ptrs, s)) return rc;
  char* flat = r.base[r.rank] + flat_off;
  const int grid = pack_grid(r, nt);
  if (wire == B2DP_F32)
    pack_kernel<float, DIR><<<grid, kPackThreads, 0, s>>>(table.abs + t0, nt, (float*)flat, prescale, skip_lo, skip_hi);
  else
    pack_kernel<__nv_bfloat16, DIR><<<grid, kPackThreads, 0, s>>>(table.abs + t0, nt, (__nv_bfloat16*)flat, prescale,
                                                                  skip_lo, skip_hi);
  CUDA_TRY(cudaGetLastError());
  return B2DP_OK;
}

template <typename KernelT>
int step_grid(KernelT kernel, const RankState& r, int64_t nvec, int U, int grid_cap = kMaxBlocks, size_t dyn_smem = 0) {
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kStepThreads, dyn_smem) != cudaSuccess || occ < 1) {
    (void)cudaGetLastError();
    occ = 1;
  }
  int64_t want = (nvec + (int64_t)kStepThreads * U - 1) / ((int64_t)kStepThreads * U);
  int64_t cap = std::min<int64_t>(std::max(1, std::min(kMaxBlocks, grid_cap)), (int64_t)r.num_sms * occ);
  if (want <= cap) return (int)std::max<int64_t>(1, want);
  // A persistent grid-stride loop runs ceil(want / grid) rounds and the last round costs a full
  // round however few blocks still have work in it.  Take the smallest grid that needs the same
  // number of rounds as the cap would: every block then does the same number of rounds (+-0)
  // and no round is spent on a fraction of the grid (up to one round in ~11 at n = 2).
  const int64_t rounds = (want + cap - 1) / cap;
  return (int)((want + rounds - 1) / rounds);
}

struct Tuning {
  int grid_cap, unroll, vec = 4;
  bool own = false;   // the rank's own slice is read in place from the gradient tensors (a.own_* set)
};

template <int N, int OPT>
int launch_step256(const RankState& r, const StepArgs& a, const Tuning& t, cudaStream_t s) {
  // N == 2: two 256-bit vectors per peer in flight per thread unless the caller asked for one
  // (r02 sweep on 2 x B200: 171.7 vs 175.6 us; with N >= 4 the second vector spills registers)
  const size_t smem = t.own ? (size_t)a.own_ntiles * kOwnTileBytes + 16 : 0;   // the bucket's tile table (own slice read in place)
  auto go = [&](auto kernel, int U) {
    const int grid = step_grid(kernel, r, a.nvec >> 1, U, t.grid_cap, smem);
    kernel<<<grid, kStepThreads, smem, s>>>(a);
    return cudaGetLastError();
  };
  cudaError_t e;
  constexpr int U2 = (N == 2 ? 2 : 1);
  const bool two = (t.unroll == 2 || t.unroll == 0) && N == 2;
  if (t.own && OPT != B2DP_OPT_NONE)
    e = two ? go(step256_kernel<N, (OPT == B2DP_OPT_NONE ? B2DP_OPT_ADAM : OPT), U2, true>, U2)
            : go(step256_kernel<N, (OPT == B2DP_OPT_NONE ? B2DP_OPT_ADAM : OPT), 1, true>, 1);
  else
    e = two ? go(step256_kernel<N, OPT, U2, false>, U2) : go(step256_kernel<N, OPT, 1, false>, 1);
  if (e != cudaSuccess) return fail(B2DP_ERR_CUDA, "step launch: %s", cudaGetErrorString(e));
  return B2DP_OK;
}

template <int OPT>
int launch_step256_n(int n, const RankState& r, const StepArgs& a, const Tuning& t, cudaStream_t s) {
  switch (n) {
    case 2: return launch_step256<2, OPT>(r, a, t, s);
    case 4: return launch_step256<4, OPT>(r, a, t, s);
    case 8: return launch_step256<8, OPT>(r, a, t, s);
  }
  return fail(B2DP_ERR_INVALID, "256-bit peer kernel is built for 2, 4 and 8 ranks (got %d)", n);
}

template <int N, typename WireT, int OPT, int U>
int launch_step_u(const RankState& r, const StepArgs& a, const Tuning& t, cudaStream_t s) {
  auto kernel = step_kernel<N, WireT, OPT, U>;
  const int grid = step_grid(kernel, r, a.nvec, U, t.grid_cap);
  kernel<<<grid, kStepThreads, 0, s>>>(a);
  CUDA_TRY(cudaGetLastError());
  return B2DP_OK;
}

template <int N, typename WireT, int OPT>
int launch_step_n(const RankState& r, const StepArgs& a, const Tuning& t, cudaStream_t s) {
  const int u = t.unroll ? t.unroll : ((N >= 8) ? 1 : 2);
  if (u == 1) return launch_step_u<N, WireT, OPT, 1>(r, a, t, s);
  return launch_step_u<N, WireT, OPT, 2>(r, a, t, s);
}

template <typename WireT, int OPT>
int launch_step_w(int n, const RankState& r, const StepArgs& a, const Tuning& t, cudaStream_t s) {
  switch (n) {
    case 1: return launch_step_n<1, WireT, OPT>(r, a, t, s);
    case 2: return launch_step_n<2, WireT, OPT>(r, a, t, s);
    case 3: return launch_step_n<3, WireT, OPT>(r, a, t, s);
    case 4: return launch_step_n<4, WireT, OPT>(r, a, t, s);
    case 5: return launch_step_n<5, WireT, OPT>(r, a, t, s);
    case 6: return launch_step_n<6, WireT, OPT>(r, a, t, s);
    case 7: return launch_step_n<7, WireT, OPT>(r, a, t, s);
    case 8: return launch_step_n<8, WireT, OPT>(r, a, t, s);
  }
  return fail(B2DP_ERR_INVALID, "unsupported rank count %d", n);
}

template <int OPT>
int launch_step_o(int wire, int n, const RankState& r, const StepArgs& a, const Tuning& t, cudaStream_t s) {
  if (wire == B2DP_F32 && t.vec == 8 && (n == 2 || n == 4 || n == 8)) return launch_step256_n<OPT>(n, r, a, t, s);
  if (wire == B2DP_F32) return launch_step_w<float, OPT>(n, r, a, t, s);
  return launch_step_w<__nv_bfloat16, OPT>(n, r, a, t, s);
}

int launch_gather(int n, const RankState& r, const StepArgs& a, cudaStream_t s) {
  auto go = [&](auto kernel) {
    const int grid = step_grid(kernel, r, a.nvec, 1);
    kernel<<<grid, kStepThreads, 0, s>>>(a);
    return cudaGetLastError();
  };
  cudaError_t e = cudaSuccess;
  switch (n) {
    case 1: return B2DP_OK;
    case 2: e = go(gather_kernel<2>); break;
    case 3: e = go(gather_kernel<3>); break;
    case 4: e = go(gather_kernel<4>); break;
    case 5: e = go(gather_kernel<5>); break;
    case 6: e = go(gather_kernel<6>); break;
    case 7: e = go(gather_kernel<7>); break;
    case 8: e = go(gather_kernel<8>); break;
    default: return fail(B2DP_ERR_INVALID, "unsupported rank count %d", n);
  }
  if (e != cudaSuccess) return fail(B2DP_ERR_CUDA, "gather launch: %s", cudaGetErrorString(e));
  return B2DP_OK;
}

template <int OPT, int U>
int launch_step_mc_u(const RankState& r, const StepArgs& a, const Tuning& t, cudaStream_t s) {
  auto kernel = step_mc_kernel<OPT, U>;
  const int grid = step_grid(kernel, r, a.nvec, U, t.grid_cap);
  kernel<<<grid, kStepThreads, 0, s>>>(a);
  CUDA_TRY(cudaGetLastError());
  return B2DP_OK;
}

template <int OPT>
int launch_step_mc_o(const RankState& r, const StepArgs& a, const Tuning& t, cudaStream_t s) {
  if (t.unroll == 1) return launch_step_mc_u<OPT, 1>(r, a, t, s);
  if (t.unroll == 2) return launch_step_mc_u<OPT, 2>(r, a, t, s);
  return launch_step_mc_u<OPT, 4>(r, a, t, s);
}

int launch_step_mc(int opt, const RankState& r, const StepArgs& a, const Tuning& t, cudaStream_t s) {
  if (opt == B2DP_OPT_ADAM) return launch_step_mc_o<B2DP_OPT_ADAM>(r, a, t, s);
  if (opt == B2DP_OPT_SGDM) return launch_step_mc_o<B2DP_OPT_SGDM>(r, a, t, s);
  return launch_step_mc_o<B2DP_OPT_NONE>(r, a, t, s);
}

int launch_step_mc_bf16(int opt, const RankState& r, const StepArgs& a, const Tuning& t, cudaStream_t s) {
  auto go = [&](auto kernel) {
    const int grid = step_grid(kernel, r, a.nvec >> 1, 1, t.grid_cap);
    kernel<<<grid, kStepThreads, 0, s>>>(a);
    return cudaGetLastError();
  };
  cudaError_t e;
  if (opt == B2DP_OPT_ADAM)
    e = go(step_mc_bf16_kernel<B2DP_OPT_ADAM>);
  else if (opt == B2DP_OPT_SGDM)
    e = go(step_mc_bf16_kernel<B2DP_OPT_SGDM>);
  else
    e = go(step_mc_bf16_kernel<B2DP_OPT_NONE>);
  if (e != cudaSuccess) return fail(B2DP_ERR_CUDA, "multicast bf16 step launch: %s", cudaGetErrorString(e));
  return B2DP_OK;
}

int fill_comm(b2dp_plan* p, RankState& r, Comm* c) {
  const int n = p->ctx->nranks;
  for (int i = 0; i < kMaxRanks; ++i) c->flags[i] = nullptr;
  for (int i = 0; i < n; ++i) c->flags[i] = reinterpret_cast<uint32_t*>(r.base[i] + p->map.flags);
  c->status = reinterpret_cast<int32_t*>(r.base[r.rank] + p->map.ctrl + kCtrlStatus);
  c->host_status = r.host_status_dev;
  c->timeout_ns = (uint64_t)(p->ctx->timeout_s * 1e9);
  c->rank = r.rank;
  c->n = n;
  c->use_flags = (n > 1 && !p->ctx->emulated) ? 1 : 0;
  c->epoch = ++r.epoch;
  return B2DP_OK;
}

int fill_opt(int opt, const b2dp_hyper* h, float postscale, OptScalars* o) {
  memset(o, 0, sizeof *o);
  o->postscale = postscale;
  if (opt == B2DP_OPT_NONE) return B2DP_OK;
  if (!h) return fail(B2DP_ERR_INVALID, "hyper is null");
  o->postscale = h->postscale;
  if (opt == B2DP_OPT_ADAM) {
    if (h->step < 1) return fail(B2DP_ERR_INVALID, "Adam step must be >= 1 (got %lld)", (long long)h->step);
    o->alpha = b2dp_adam_alpha(h->lr, h->beta1, h->beta2, h->step);
    o->omb1 = 1.0f - h->beta1;
    o->omb2 = 1.0f - h->beta2;
    o->eps = h->eps;
  } else if (opt == B2DP_OPT_SGDM) {
    o->lr = h->lr;
    o->mu = h->momentum;
    o->wd = h->weight_decay;
    o->nesterov = h->nesterov;
  } else {
    return fail(B2DP_ERR_INVALID, "unknown optimizer %d", opt);
  }
  return B2DP_OK;
}

// mode 0: fused step (push), 1: reduce-scatter + apply only, 2: all-reduce, 3: all-gather only
int run_collective(b2dp_plan* p, int dev, int bucket, int opt, const b2dp_hyper* h, float postscale, int mode,
                   void* stream, bool stats_continue = false, const float* own_prescale = nullptr) {
  Range nvtx_range("b2dp::collective");
  RankState* rp = nullptr;
  if (int rc = get_rank(p, dev, &rp)) return rc;
  RankState& r = *rp;
  if (!b2dp_plan_ready(p)) return fail(B2DP_ERR_STATE, "plan has unattached peers");
  if (int rc = check_poisoned(r)) return rc;
  int b0 = 0, b1 = 0;
  if (int rc = bucket_range(p, bucket, &b0, &b1)) return rc;
  OptScalars o;
  if (int rc = fill_opt(opt, h, postscale, &o)) return rc;
  bool stats_started = stats_continue;  // the first bucket launched by a call stores STATS, later ones add
  const int n = p->ctx->nranks;
  const size_t wire = p->layout.wire == B2DP_BF16 ? 2 : 4;
  cudaStream_t s = (cudaStream_t)stream;
  DeviceGuard guard(r.device);
  for (int b = b0; b < b1; ++b) {
    const Bucket& bk = p->layout.buckets[b];
    if (bk.slice == 0) continue;
    StepArgs a;
    memset(&a, 0, sizeof a);
    const size_t slice_elem0 = (size_t)bk.offset + (size_t)r.rank * bk.slice;
    const size_t out_off = (mode == 2) ? p->map.reduced : p->map.weights;
    for (int i = 0; i < n; ++i) {
      a.g[i] = r.base[i] + p->map.grads + slice_elem0 * wire;
      a.out[i] = reinterpret_cast<float*>(r.base[i] + out_off + slice_elem0 * 4);
    }
    a.w = a.out[r.rank];
    a.s0 = reinterpret_cast<float*>(r.base[r.rank] + p->map.s0 + (size_t)bk.shard_off * 4);
    a.s1 = reinterpret_cast<float*>(r.base[r.rank] + p->map.s1 + (size_t)bk.shard_off * 4);
    a.nvec = bk.slice / 4;
    a.o = o;
    a.partials = reinterpret_cast<double*>(r.base[r.rank] + p->map.partials);
    a.ticket = reinterpret_cast<uint32_t*>(r.base[r.rank] + p->map.ctrl + kCtrlTicket);
    a.stats = reinterpret_cast<double*>(r.base[r.rank] + p->map.ctrl + kCtrlStats);
    a.want_stats = (h && h->want_stats) ? 1 : 0;
    a.stats_add = stats_started ? 1 : 0;
    stats_started = true;
    a.push = (mode == 1) ? 0 : 1;
    fill_comm(p, r, &a.c);
    Tuning peer_tune{p->grid_cap, p->peer_unroll, p->peer_vec};
    if (own_prescale && bk.tile_count <= kOwnMaxTiles) {   // b2dp_pack_step left this rank's own slice unpacked: the kernel reads it in place
      a.own_tiles = r.grads.abs + bk.tile_first;
      a.own_ntiles = (int32_t)bk.tile_count;
      a.own_bucket_off = bk.offset;
      a.own_rel8 = ((int64_t)r.rank * bk.slice) >> 3;
      a.prescale = *own_prescale;
      peer_tune.own = true;
    }
    int rc;
    if (n == 1 && p->layout.wire == B2DP_F32 && opt != B2DP_OPT_NONE && mode != 3) {
      const int64_t nvec8 = a.nvec >> 1;
      const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((nvec8 + kApplyThreads - 1) / kApplyThreads, kApplyMaxBlocks));
      if (opt == B2DP_OPT_ADAM)
        apply1_kernel<B2DP_OPT_ADAM><<<grid, kApplyThreads, 0, s>>>(a);
      else
        apply1_kernel<B2DP_OPT_SGDM><<<grid, kApplyThreads, 0, s>>>(a);
      CUDA_TRY(cudaGetLastError());
      continue;
    }
    const bool use_mc = p->path == B2DP_PATH_MULTICAST && a.c.use_flags;
    if (use_mc && mode == 3) {
      char* mc = reinterpret_cast<char*>(r.mc_va);
      a.out[0] = reinterpret_cast<float*>(mc + out_off + slice_elem0 * 4);
      const int grid = step_grid(gather_mc_kernel, r, a.nvec, 1, p->mc_grid_cap);
      gather_mc_kernel<<<grid, kStepThreads, 0, s>>>(a);
      CUDA_TRY(cudaGetLastError());
      rc = B2DP_OK;
    } else if (use_mc) {
      char* mc = reinterpret_cast<char*>(r.mc_va);
      a.g[0] = mc + p->map.grads + slice_elem0 * wire;
      a.out[0] = reinterpret_cast<float*>(mc + out_off + slice_elem0 * 4);
      a.w = reinterpret_cast<float*>(r.base[r.rank] + out_off + slice_elem0 * 4);
      if (p->layout.wire == B2DP_BF16)
        rc = launch_step_mc_bf16(opt, r, a, Tuning{p->mc_bf16_grid_cap, 1}, s);
      else
        rc = launch_step_mc(opt, r, a, Tuning{p->mc_grid_cap, p->mc_unroll}, s);
    } else if (mode == 3)
      rc = launch_gather(n, r, a, s);
    else if (opt == B2DP_OPT_ADAM)
      rc = launch_step_o<B2DP_OPT_ADAM>(p->layout.wire, n, r, a, peer_tune, s);
    else if (opt == B2DP_OPT_SGDM)
      rc = launch_step_o<B2DP_OPT_SGDM>(p->layout.wire, n, r, a, peer_tune, s);
    else
      rc = launch_step_o<B2DP_OPT_NONE>(p->layout.wire, n, r, a, peer_tune, s);
    if (rc) return rc;
  }
  return B2DP_OK;
}

void destroy_rank(RankState& r, size_t block_bytes = 0, bool local = true) {
  if (r.device < 0) return;
  DeviceGuard guard(r.device);
  cudaDeviceSynchronize();
  if (r.d_tiles) cudaFree(r.d_tiles);
  for (PtrTable* t : {&r.grads, &r.outs, &r.weights}) {
    if (t->dev) cudaFree(t->dev);
    if (t->abs) cudaFree(t->abs);
    t->dev = nullptr;
    t->abs = nullptr;
  }
  if (r.host_status) cudaFreeHost(r.host_status);
  r.host_status = r.host_status_dev = nullptr;
  if (r.side) cudaStreamDestroy(r.side);
  if (r.ev_start) cudaEventDestroy(r.ev_start);
  for (cudaEvent_t e : r.ev_packed) cudaEventDestroy(e);
  r.side = nullptr;
  r.ev_start = nullptr;
  r.ev_packed.clear();
  // local context: every rank shares one multicast handle/mapping and the owner frees it;
  // rank context: this process holds its own (created or imported) handle and mapping
  if (r.mc_handle && (r.mc_owner || !local)) mc_release(r.mc_handle, r.mc_va, block_bytes);
  for (auto& b : r.imported) vmm_free(&b);
  vmm_free(&r.own);
  (void)cudaGetLastError();
}

}  // namespace

// ------------------------------------------------------------------------------ context
extern "C" int b2dp_ctx_create_local(int32_t ndev, const int32_t* devices, b2dp_ctx** out) {
  if (!out) return fail(B2DP_ERR_INVALID, "ctx_create_local: out is null");
  *out = nullptr;
  if (ndev < 1 || ndev > kMaxRanks || !devices)
    return fail(B2DP_ERR_INVALID, "ctx_create_local: ndev %d outside 1..%d", ndev, kMaxRanks);
  auto c = std::make_unique<b2dp_ctx>();
  c->local = true;
  c->nranks = c->nlocal = ndev;
  for (int i = 0; i < ndev; ++i) {
    if (int rc = check_device(devices[i])) return rc;
    c->devices[i] = devices[i];
    c->ranks[i] = i;
    for (int j = 0; j < i; ++j) c->emulated |= (devices[j] == devices[i]);
  }
  if (int rc = vmm_init()) return rc;
  *out = c.release();
  return B2DP_OK;
}

extern "C" int b2dp_ctx_create_rank(int32_t rank, int32_t nranks, int32_t device, b2dp_ctx** out) {
  if (!out) return fail(B2DP_ERR_INVALID, "ctx_create_rank: out is null");
  *out = nullptr;
  if (nranks < 1 || nranks > kMaxRanks || rank < 0 || rank >= nranks)
    return fail(B2DP_ERR_INVALID, "ctx_create_rank: rank %d of %d (at most %d ranks)", rank, nranks, kMaxRanks);
  if (int rc = check_device(device)) return rc;
  if (int rc = vmm_init()) return rc;
  auto c = std::make_unique<b2dp_ctx>();
  c->local = false;
  c->nranks = nranks;
  c->nlocal = 1;
  c->devices[0] = device;
  c->ranks[0] = rank;
  *out = c.release();
  return B2DP_OK;
}

extern "C" void b2dp_ctx_destroy(b2dp_ctx* c) { delete c; }
extern "C" int32_t b2dp_ctx_nranks(const b2dp_ctx* c) { return c ? c->nranks : -1; }
extern "C" int32_t b2dp_ctx_num_local(const b2dp_ctx* c) { return c ? c->nlocal : -1; }
extern "C" int b2dp_ctx_set_timeout(b2dp_ctx* c, double seconds) {
  if (!c || !(seconds > 0)) return fail(B2DP_ERR_INVALID, "set_timeout: need a context and seconds > 0");
  c->timeout_s = seconds;
  return B2DP_OK;
}

// --------------------------------------------------------------------------------- plan
namespace {
// Builds one local rank of a plan; on any failure everything this rank already holds is
// released before returning, so b2dp_plan_create never leaks a half-built rank.
int build_rank(b2dp_plan* p, const b2dp_ctx* c, const b2dp_layout* l, int i, std::unique_ptr<RankState>* out) {
  auto r = std::make_unique<RankState>();
  r->device = c->devices[i];
  r->rank = c->ranks[i];
  DeviceGuard guard(r->device);
  auto body = [&]() -> int {
    CUDA_TRY(cudaFree(0));  // make sure the primary context exists before driver calls
    CUDA_TRY(cudaDeviceGetAttribute(&r->num_sms, cudaDevAttrMultiProcessorCount, r->device));
    size_t gran = 0, mgran = 0;
    if (int rc = vmm_granularity(r->device, &gran)) return rc;
    if (c->nranks > 1 && !c->emulated && mc_granularity(c->nranks, 1 << 21, &mgran) == B2DP_OK && mgran > gran)
      gran = mgran;  // keep the block bindable to a multicast object
    p->block_bytes = align_up(p->map.end, gran);
    if (int rc = vmm_alloc(r->device, p->block_bytes, c->devices, c->nlocal, &r->own)) return rc;
    CUDA_TRY(cudaMemset(reinterpret_cast<void*>(r->own.va), 0, p->map.end));
    if (!l->tiles.empty()) {
      CUDA_TRY(cudaMalloc(&r->d_tiles, sizeof(Tile) * l->tiles.size()));
      CUDA_TRY(cudaMemcpy(r->d_tiles, l->tiles.data(), sizeof(Tile) * l->tiles.size(), cudaMemcpyHostToDevice));
      for (PtrTable* t : {&r->grads, &r->outs, &r->weights}) {   // everything the hot path will ever need
        CUDA_TRY(cudaMalloc(&t->abs, sizeof(TileAbs) * l->tiles.size()));
        if (l->K > kParamPtrs) CUDA_TRY(cudaMalloc(&t->dev, sizeof(void*) * l->K));
      }
    }
    if (c->nranks > 1 && !c->emulated && l->buckets.size() > 1) {
      CUDA_TRY(cudaStreamCreateWithFlags(&r->side, cudaStreamNonBlocking));
      CUDA_TRY(cudaEventCreateWithFlags(&r->ev_start, cudaEventDisableTiming));
      for (size_t b = 0; b < l->buckets.size(); ++b) {
        cudaEvent_t e = nullptr;
        CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        r->ev_packed.push_back(e);
      }
    }
    if (c->nranks > 1 && !c->emulated) {
      CUDA_TRY(cudaHostAlloc(reinterpret_cast<void**>(&r->host_status), sizeof(int32_t), cudaHostAllocMapped));
      *r->host_status = 0;
      CUDA_TRY(cudaHostGetDevicePointer(reinterpret_cast<void**>(&r->host_status_dev), r->host_status, 0));
    }
    return B2DP_OK;
  };
  if (int rc = body()) {
    destroy_rank(*r);
    return rc;
  }
  r->base[r->rank] = reinterpret_cast<char*>(r->own.va);
  r->attached[r->rank] = true;
  *out = std::move(r);
  return B2DP_OK;
}
}  // namespace

extern "C" int b2dp_plan_create(b2dp_ctx* c, const b2dp_layout* l, b2dp_plan** out) {
  if (!out) return fail(B2DP_ERR_INVALID, "plan_create: out is null");
  *out = nullptr;
  if (!c || !l) return fail(B2DP_ERR_INVALID, "plan_create: null context or layout");
  if (l->nranks != c->nranks)
    return fail(B2DP_ERR_INVALID, "plan_create: layout is cut for %d ranks, context has %d", l->nranks, c->nranks);
  auto p = std::make_unique<b2dp_plan>();
  p->ctx = c;
  p->ctx_local = c->local;
  p->layout = *l;
  p->map = make_map(*l);
  for (int i = 0; i < c->nlocal; ++i) {
    std::unique_ptr<RankState> r;
    if (int rc = build_rank(p.get(), c, l, i, &r)) {
      for (auto& q : p->rs) destroy_rank(*q);  // ranks finished earlier; the failed one cleaned up after itself
      return rc;
    }
    p->rs.push_back(std::move(r));
  }
  if (c->local) {  // one address space: every rank sees every block at the same address
    for (auto& r : p->rs)
      for (auto& q : p->rs) {
        r->base[q->rank] = reinterpret_cast<char*>(q->own.va);
        r->attached[q->rank] = true;
      }
  }
  *out = p.release();
  return B2DP_OK;
}

extern "C" void b2dp_plan_destroy(b2dp_plan* p) {
  if (!p) return;
  for (auto& r : p->rs) destroy_rank(*r, p->block_bytes, p->ctx_local);
  delete p;
}

extern "C" int b2dp_plan_export_fd(b2dp_plan* p, int* fd_out) {
  if (!p || !fd_out) return fail(B2DP_ERR_INVALID, "export_fd: null argument");
  if (p->ctx->local) return fail(B2DP_ERR_STATE, "export_fd: a local context maps its peers itself");
  return vmm_export_fd(p->rs[0]->own, fd_out);
}

extern "C" int b2dp_plan_attach_peer(b2dp_plan* p, int32_t peer, int fd) {
  if (!p) return fail(B2DP_ERR_INVALID, "attach_peer: plan is null");
  if (p->ctx->local) return fail(B2DP_ERR_STATE, "attach_peer: a local context maps its peers itself");
  RankState& r = *p->rs[0];
  if (peer < 0 || peer >= p->ctx->nranks) return fail(B2DP_ERR_INVALID, "attach_peer: rank %d out of range", peer);
  if (peer == r.rank) return B2DP_OK;
  if (r.attached[peer]) return fail(B2DP_ERR_STATE, "attach_peer: rank %d is already attached", peer);
  DeviceGuard guard(r.device);
  // map the exporter's WHOLE allocation (it was rounded to the multicast granularity, which can
  // exceed this device's own allocation granularity), not just the bytes the layout uses
  if (int rc = vmm_import_fd(fd, p->block_bytes, r.device, &r.imported[peer])) return rc;
  r.base[peer] = reinterpret_cast<char*>(r.imported[peer].va);
  r.attached[peer] = true;
  return B2DP_OK;
}

extern "C" int32_t b2dp_plan_ready(const b2dp_plan* p) {
  if (!p) return 0;
  for (auto& r : p->rs)
    for (int i = 0; i < p->ctx->nranks; ++i)
      if (!r->attached[i]) return 0;
  return 1;
}

extern "C" int b2dp_plan_buffer(b2dp_plan* p, int32_t dev, int32_t which, void** ptr, int64_t* elems) {
  RankState* r = nullptr;
  if (int rc = get_rank(p, dev, &r)) return rc;
  size_t off;
  int64_t n;
  switch (which) {
    case B2DP_BUF_GRADS: off = p->map.grads; n = p->layout.total; break;
    case B2DP_BUF_WEIGHTS: off = p->map.weights; n = p->layout.total; break;
    case B2DP_BUF_REDUCED: off = p->map.reduced; n = p->layout.total; break;
    case B2DP_BUF_STATE0: off = p->map.s0; n = p->layout.shard; break;
    case B2DP_BUF_STATE1: off = p->map.s1; n = p->layout.shard; break;
    case B2DP_BUF_STATS: off = p->map.ctrl + kCtrlStats; n = 4; break;
    default: return fail(B2DP_ERR_INVALID, "plan_buffer: unknown buffer %d", which);
  }
  if (ptr) *ptr = r->base[r->rank] + off;
  if (elems) *elems = n;
  return B2DP_OK;
}

extern "C" int b2dp_plan_check(b2dp_plan* p, int32_t dev, void* stream) {
  RankState* r = nullptr;
  if (int rc = get_rank(p, dev, &r)) return rc;
  DeviceGuard guard(r->device);
  cudaStream_t s = (cudaStream_t)stream;
  int32_t st = 0;
  void* word = r->base[r->rank] + p->map.ctrl + kCtrlStatus;
  CUDA_TRY(cudaMemcpyAsync(&st, word, sizeof st, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  if (st == 0 && r->host_status) st = *reinterpret_cast<volatile int32_t*>(r->host_status);
  if (st != 0) {
    CUDA_TRY(cudaMemsetAsync(word, 0, sizeof st, s));
    if (r->host_status) *reinterpret_cast<volatile int32_t*>(r->host_status) = 0;
    return fail(st, "rank %d: a cross-GPU flag wait timed out (a peer never reached the collective)", r->rank);
  }
  return B2DP_OK;
}

// ------------------------------------------------------------- Path B: NVSwitch multicast
extern "C" int b2dp_plan_multicast_create(b2dp_plan* p, int* fd_out) {
  if (!p) return fail(B2DP_ERR_INVALID, "multicast_create: plan is null");
  if (fd_out) *fd_out = -1;
  const int n = p->ctx->nranks;
  if (n < 2 || p->ctx->emulated) return fail(B2DP_ERR_STATE, "multicast_create: needs >= 2 distinct GPUs");
  if (!b2dp_plan_ready(p)) return fail(B2DP_ERR_STATE, "multicast_create: attach the peers first");
  RankState& r0 = *p->rs[0];
  if (r0.mc_handle) return fail(B2DP_ERR_STATE, "multicast_create: already created");
  for (auto& r : p->rs)
    if (int rc = mc_supported(r->device)) return rc;
  if (!p->ctx->local && r0.rank != 0) return fail(B2DP_ERR_STATE, "multicast_create: only rank 0 creates the object");
  DeviceGuard guard(r0.device);
  unsigned long long h = 0;
  if (int rc = mc_create(n, p->block_bytes, &h)) return rc;
  if (p->ctx->local) {
    // one process: add every device, bind every block, map once for all devices
    for (auto& r : p->rs)
      if (int rc = mc_add_device(h, r->device)) return rc;
    for (auto& r : p->rs) {
      DeviceGuard g2(r->device);
      if (int rc = mc_bind(h, r->own, p->block_bytes)) return rc;
    }
    unsigned long long va = 0;
    if (int rc = mc_map(h, p->block_bytes, p->ctx->devices, p->ctx->nlocal, &va)) return rc;
    for (auto& r : p->rs) {
      r->mc_handle = h;
      r->mc_va = va;
      r->mc_joined = r->mc_bound = true;
    }
    r0.mc_owner = true;
    return B2DP_OK;
  }
  r0.mc_handle = h;
  r0.mc_owner = true;
  if (!fd_out) return fail(B2DP_ERR_INVALID, "multicast_create: fd_out is required for a rank context");
  return mc_export_fd(h, fd_out);
}

extern "C" int b2dp_plan_multicast_join(b2dp_plan* p, int fd) {
  if (!p) return fail(B2DP_ERR_INVALID, "multicast_join: plan is null");
  if (p->ctx->local) return fail(B2DP_ERR_STATE, "multicast_join: a local context joins inside multicast_create");
  RankState& r = *p->rs[0];
  if (r.mc_joined) return fail(B2DP_ERR_STATE, "multicast_join: already joined");
  DeviceGuard guard(r.device);
  if (!r.mc_owner) {
    if (int rc = mc_supported(r.device)) return rc;
    if (int rc = mc_import_fd(fd, &r.mc_handle)) return rc;
  }
  if (int rc = mc_add_device(r.mc_handle, r.device)) return rc;
  r.mc_joined = true;
  return B2DP_OK;
}

extern "C" int b2dp_plan_multicast_bind(b2dp_plan* p) {
  if (!p) return fail(B2DP_ERR_INVALID, "multicast_bind: plan is null");
  if (p->ctx->local) return fail(B2DP_ERR_STATE, "multicast_bind: a local context binds inside multicast_create");
  RankState& r = *p->rs[0];
  if (!r.mc_joined) return fail(B2DP_ERR_STATE, "multicast_bind: join first (and wait until every rank has joined)");
  if (r.mc_bound) return B2DP_OK;
  DeviceGuard guard(r.device);
  if (int rc = mc_bind(r.mc_handle, r.own, p->block_bytes)) return rc;
  if (int rc = mc_map(r.mc_handle, p->block_bytes, &r.device, 1, &r.mc_va)) return rc;
  r.mc_bound = true;
  return B2DP_OK;
}

extern "C" int b2dp_plan_set_path(b2dp_plan* p, int32_t path) {
  if (!p) return fail(B2DP_ERR_INVALID, "set_path: plan is null");
  if (path == B2DP_PATH_PEER) {
    p->path = path;
    return B2DP_OK;
  }
  if (path != B2DP_PATH_MULTICAST) return fail(B2DP_ERR_INVALID, "set_path: unknown path %d", path);
  for (auto& r : p->rs)
    if (!r->mc_bound) return fail(B2DP_ERR_STATE, "set_path: the plan's multicast object is not bound (multicast_create/join/bind)");
  p->path = path;
  return B2DP_OK;
}

extern "C" int32_t b2dp_plan_path(const b2dp_plan* p) { return p ? p->path : -1; }

extern "C" int b2dp_plan_set_tuning(b2dp_plan* p, int32_t key, int32_t value) {
  if (!p) return fail(B2DP_ERR_INVALID, "set_tuning: plan is null");
  switch (key) {
    case B2DP_TUNE_GRID_CAP:
      if (value < 1) return fail(B2DP_ERR_INVALID, "set_tuning: grid cap must be >= 1");
      p->grid_cap = value;
      return B2DP_OK;
    case B2DP_TUNE_MC_GRID_CAP:
      if (value < 1) return fail(B2DP_ERR_INVALID, "set_tuning: grid cap must be >= 1");
      p->mc_grid_cap = value;
      return B2DP_OK;
    case B2DP_TUNE_MC_BF16_GRID_CAP:
      if (value < 1) return fail(B2DP_ERR_INVALID, "set_tuning: grid cap must be >= 1");
      p->mc_bf16_grid_cap = value;
      return B2DP_OK;
    case B2DP_TUNE_SKIP_OWN:
      if (value != 0 && value != 1) return fail(B2DP_ERR_INVALID, "set_tuning: skip_own is 0 or 1");
      p->skip_own = value;
      return B2DP_OK;
    case B2DP_TUNE_MC_UNROLL:
      if (value != 1 && value != 2 && value != 4) return fail(B2DP_ERR_INVALID, "set_tuning: multicast unroll is 1, 2 or 4");
      p->mc_unroll = value;
      return B2DP_OK;
    case B2DP_TUNE_PEER_VEC:
      if (value != 4 && value != 8) return fail(B2DP_ERR_INVALID, "set_tuning: peer vector width is 4 or 8 elements");
      p->peer_vec = value;
      return B2DP_OK;
    case B2DP_TUNE_PEER_UNROLL:
      if (value < 0 || value > 2) return fail(B2DP_ERR_INVALID, "set_tuning: peer unroll is 0 (auto), 1 or 2");
      p->peer_unroll = value;
      return B2DP_OK;
  }
  return fail(B2DP_ERR_INVALID, "set_tuning: unknown key %d", key);
}

// ------------------------------------------------------------------------- the hot path
extern "C" int b2dp_pack(b2dp_plan* p, int32_t dev, int32_t bucket, const void* const* grads,
                         int32_t grad_dtype, float prescale, void* stream) {
  Range nvtx_range("b2dp::pack");
  RankState* r = nullptr;
  if (int rc = get_rank(p, dev, &r)) return rc;
  if (grad_dtype != B2DP_F32) return fail(B2DP_ERR_INVALID, "pack: gradients must be f32 (got dtype %d)", grad_dtype);
  DeviceGuard guard(r->device);
  return launch_pack<0>(p, *r, bucket, r->grads, grads, p->map.grads, p->layout.wire, prescale,
                        (cudaStream_t)stream);
}

extern "C" int b2dp_load_weights(b2dp_plan* p, int32_t dev, const void* const* weights, void* stream) {
  RankState* r = nullptr;
  if (int rc = get_rank(p, dev, &r)) return rc;
  DeviceGuard guard(r->device);
  return launch_pack<0>(p, *r, -1, r->weights, weights, p->map.weights, B2DP_F32, 1.0f, (cudaStream_t)stream);
}

extern "C" int b2dp_unpack(b2dp_plan* p, int32_t dev, int32_t bucket, int32_t which, void* const* outs,
                           void* stream) {
  Range nvtx_range("b2dp::unpack");
  RankState* r = nullptr;
  if (int rc = get_rank(p, dev, &r)) return rc;
  size_t off;
  if (which == B2DP_BUF_WEIGHTS)
    off = p->map.weights;
  else if (which == B2DP_BUF_REDUCED)
    off = p->map.reduced;
  else
    return fail(B2DP_ERR_INVALID, "unpack: buffer %d is not WEIGHTS or REDUCED", which);
  DeviceGuard guard(r->device);
  return launch_pack<1>(p, *r, bucket, r->outs, (const void* const*)outs, off, B2DP_F32, 1.0f,
                        (cudaStream_t)stream);
}

// pack + step in one call.  n == 1 with an f32 wire: ONE kernel that reads the gradients in
// place (no bucket copy at all); otherwise pack followed by the fused collective step.
extern "C" int b2dp_pack_step(b2dp_plan* p, int32_t dev, int32_t bucket, const void* const* grads, int32_t grad_dtype,
                              float prescale, int32_t opt, const b2dp_hyper* hyper, void* stream) {
  Range nvtx_range("b2dp::pack_step");
  RankState* rp = nullptr;
  if (int rc = get_rank(p, dev, &rp)) return rc;
  if (opt != B2DP_OPT_ADAM && opt != B2DP_OPT_SGDM) return fail(B2DP_ERR_INVALID, "pack_step: optimizer must be ADAM or SGDM");
  if (grad_dtype != B2DP_F32) return fail(B2DP_ERR_INVALID, "pack_step: gradients must be f32 (got dtype %d)", grad_dtype);
  int b0 = 0, b1 = 0;
  if (int rc = bucket_range(p, bucket, &b0, &b1)) return rc;
  int64_t nt = 0;
  for (int b = b0; b < b1; ++b) nt += p->layout.buckets[b].tile_count;
  const bool direct = p->ctx->nranks == 1 && p->layout.wire == B2DP_F32 &&
                      (!(hyper && hyper->want_stats) || nt * kSubTiles <= kApplyMaxBlocks);
  if (!direct) {
    RankState& r = *rp;
    const int n = p->ctx->nranks;
    // Peer path, f32 wire, 256-bit kernel, B2DP_TUNE_SKIP_OWN: rank r's own slice of a bucket is
    // read by nobody but rank r itself, so it is not packed at all -- the step kernel reads it
    // where the framework left it (1/n less pack traffic).  Same bits as the packed form.  Off by
    // default: on 2 x B200 the step kernel lost in the table walk and the cold read what the pack
    // saved (203.7-210.7 us per call against 205-206 us).
    const bool own = p->skip_own && n > 1 && !p->ctx->emulated && p->path == B2DP_PATH_PEER &&
                     p->layout.wire == B2DP_F32 && p->peer_vec == 8 && (n == 2 || n == 4 || n == 8);
    auto pack_bucket = [&](int b, cudaStream_t ps) -> int {
      int64_t lo = 0, hi = 0;
      if (own && p->layout.buckets[b].tile_count <= kOwnMaxTiles) {   // same condition as in run_collective
        const Bucket& bk = p->layout.buckets[b];
        lo = bk.offset + (int64_t)r.rank * bk.slice;
        hi = lo + bk.slice;
      }
      return launch_pack<0>(p, r, b, r.grads, grads, p->map.grads, p->layout.wire, prescale, ps, lo, hi);
    };
    const float* own_prescale = own ? &prescale : nullptr;
    cudaStream_t s = (cudaStream_t)stream;
    DeviceGuard guard(r.device);
    if (r.side && b1 - b0 > 1) {
      // Several buckets, real ranks: the step of bucket b is NVLink-bound and leaves HBM idle, so
      // bucket b+1 is packed underneath it on a side stream.  Same kernels, same arguments, same
      // order of collectives on the caller's stream as the serial form: bit-identical results.
      CUDA_TRY(cudaEventRecord(r.ev_start, s));          // gradients exist; earlier steps have read the buckets
      CUDA_TRY(cudaStreamWaitEvent(r.side, r.ev_start, 0));
      bool stats_on = false;
      for (int b = b0; b < b1; ++b) {
        if (int rc = pack_bucket(b, r.side)) return rc;
        CUDA_TRY(cudaEventRecord(r.ev_packed[b], r.side));
        CUDA_TRY(cudaStreamWaitEvent(s, r.ev_packed[b], 0));
        if (int rc = run_collective(p, dev, b, opt, hyper, 1.0f, 0, stream, stats_on, own_prescale)) return rc;
        stats_on = true;
      }
      return B2DP_OK;
    }
    if (own) {
      bool stats_on = false;
      for (int b = b0; b < b1; ++b) {
        if (int rc = pack_bucket(b, s)) return rc;
        if (int rc = run_collective(p, dev, b, opt, hyper, 1.0f, 0, stream, stats_on, own_prescale)) return rc;
        stats_on = true;
      }
      return B2DP_OK;
    }
    if (int rc = b2dp_pack(p, dev, bucket, grads, grad_dtype, prescale, stream)) return rc;
    return b2dp_step(p, dev, bucket, opt, hyper, stream);
  }
  RankState& r = *rp;
  OptScalars o;
  if (int rc = fill_opt(opt, hyper, 1.0f, &o)) return rc;
  if (nt == 0) return B2DP_OK;
  if (!grads) return fail(B2DP_ERR_INVALID, "pack_step: tensor pointer array is null");
  for (int b = b0; b < b1; ++b) {
    const Bucket& bk = p->layout.buckets[b];
    for (int k = bk.first; k < bk.first + bk.count; ++k)
      if (p->layout.numel[k] > 0 && !grads[k]) return fail(B2DP_ERR_INVALID, "tensor %d pointer is null", k);
  }
  cudaStream_t s = (cudaStream_t)stream;
  DeviceGuard guard(r.device);
  if (int rc = sync_tiles(p->layout, r.d_tiles, r.grads, grads, s)) return rc;
  StepArgs a;
  memset(&a, 0, sizeof a);
  // tile offsets are flat element offsets; with one rank a shard offset equals the flat offset
  a.w = reinterpret_cast<float*>(r.base[r.rank] + p->map.weights);
  a.s0 = reinterpret_cast<float*>(r.base[r.rank] + p->map.s0);
  a.s1 = reinterpret_cast<float*>(r.base[r.rank] + p->map.s1);
  a.o = o;
  a.partials = reinterpret_cast<double*>(r.base[r.rank] + p->map.partials);
  a.ticket = reinterpret_cast<uint32_t*>(r.base[r.rank] + p->map.ctrl + kCtrlTicket);
  a.stats = reinterpret_cast<double*>(r.base[r.rank] + p->map.ctrl + kCtrlStats);
  a.want_stats = hyper->want_stats ? 1 : 0;
  const TileAbs* tiles = r.grads.abs + p->layout.buckets[b0].tile_first;
  const unsigned grid = (unsigned)(nt * kSubTiles);
  if (opt == B2DP_OPT_ADAM)
    fused_tiles_kernel<B2DP_OPT_ADAM><<<grid, kApplyThreads, 0, s>>>(tiles, prescale, a);
  else
    fused_tiles_kernel<B2DP_OPT_SGDM><<<grid, kApplyThreads, 0, s>>>(tiles, prescale, a);
  CUDA_TRY(cudaGetLastError());
  return B2DP_OK;
}

extern "C" int b2dp_step(b2dp_plan* p, int32_t dev, int32_t bucket, int32_t opt, const b2dp_hyper* hyper,
                         void* stream) {
  if (opt == B2DP_OPT_NONE) return fail(B2DP_ERR_INVALID, "step: optimizer NONE -- use b2dp_allreduce");
  return run_collective(p, dev, bucket, opt, hyper, 1.0f, 0, stream);
}

extern "C" int b2dp_reduce_scatter_apply(b2dp_plan* p, int32_t dev, int32_t bucket, int32_t opt,
                                         const b2dp_hyper* hyper, void* stream) {
  if (opt == B2DP_OPT_NONE) return fail(B2DP_ERR_INVALID, "reduce_scatter_apply: optimizer NONE -- use b2dp_allreduce");
  return run_collective(p, dev, bucket, opt, hyper, 1.0f, 1, stream);
}

extern "C" int b2dp_all_gather(b2dp_plan* p, int32_t dev, int32_t bucket, void* stream) {
  return run_collective(p, dev, bucket, B2DP_OPT_NONE, nullptr, 1.0f, 3, stream);
}

extern "C" int b2dp_allreduce(b2dp_plan* p, int32_t dev, int32_t bucket, float postscale, void* stream) {
  return run_collective(p, dev, bucket, B2DP_OPT_NONE, nullptr, postscale, 2, stream);
}

// Sharded optimizer state -> full flat buffer on every rank: each rank
// pushes its shard of Adam m / v (or the SGD accumulator) into slice `rank` of every rank's
// B2DP_BUF_REDUCED; b2dp_unpack(B2DP_BUF_REDUCED) then hands it out per variable, which is
// what model.save() (imagenet-resnet50-mirror.py:83-86) needs from a sharded optimizer.
extern "C" int b2dp_gather_state(b2dp_plan* p, int32_t dev, int32_t which, void* stream) {
  Range nvtx_range("b2dp::gather_state");
  RankState* rp = nullptr;
  if (int rc = get_rank(p, dev, &rp)) return rc;
  RankState& r = *rp;
  size_t src_off;
  if (which == B2DP_BUF_STATE0)
    src_off = p->map.s0;
  else if (which == B2DP_BUF_STATE1)
    src_off = p->map.s1;
  else
    return fail(B2DP_ERR_INVALID, "gather_state: buffer %d is not STATE0 or STATE1", which);
  if (!b2dp_plan_ready(p)) return fail(B2DP_ERR_STATE, "plan has unattached peers");
  if (int rc = check_poisoned(r)) return rc;
  const int n = p->ctx->nranks;
  DeviceGuard guard(r.device);
  cudaStream_t s = (cudaStream_t)stream;
  for (const Bucket& bk : p->layout.buckets) {
    if (bk.slice == 0) continue;
    const size_t slice_elem0 = (size_t)bk.offset + (size_t)r.rank * bk.slice;
    float* src = reinterpret_cast<float*>(r.base[r.rank] + src_off + (size_t)bk.shard_off * 4);
    if (n == 1) {  // one rank: the shard is the whole bucket
      CUDA_TRY(cudaMemcpyAsync(r.base[r.rank] + p->map.reduced + slice_elem0 * 4, src, (size_t)bk.slice * 4,
                               cudaMemcpyDeviceToDevice, s));
      continue;
    }
    StepArgs a;
    memset(&a, 0, sizeof a);
    for (int i = 0; i < n; ++i) a.out[i] = reinterpret_cast<float*>(r.base[i] + p->map.reduced + slice_elem0 * 4);
    a.w = src;
    a.nvec = bk.slice / 4;
    a.push = 1;
    fill_comm(p, r, &a.c);
    a.c.rank = r.rank;
    // gather_kernel skips r == rank, so the local copy is done explicitly
    CUDA_TRY(cudaMemcpyAsync(a.out[r.rank], src, (size_t)bk.slice * 4, cudaMemcpyDeviceToDevice, s));
    if (int rc = launch_gather(n, r, a, s)) return rc;
  }
  return B2DP_OK;
}

extern "C" int b2dp_broadcast(b2dp_plan* p, int32_t dev, int32_t which, int32_t root, void* stream) {
  Range nvtx_range("b2dp::broadcast");
  RankState* rp = nullptr;
  if (int rc = get_rank(p, dev, &rp)) return rc;
  RankState& r = *rp;
  const int n = p->ctx->nranks;
  if (root < 0 || root >= n) return fail(B2DP_ERR_INVALID, "broadcast: root %d outside 0..%d", root, n - 1);
  size_t off;
  if (which == B2DP_BUF_WEIGHTS)
    off = p->map.weights;
  else if (which == B2DP_BUF_REDUCED)
    off = p->map.reduced;
  else
    return fail(B2DP_ERR_INVALID, "broadcast: buffer %d is not WEIGHTS or REDUCED", which);
  if (!b2dp_plan_ready(p)) return fail(B2DP_ERR_STATE, "plan has unattached peers");
  if (int rc = check_poisoned(r)) return rc;
  if (n == 1 || p->layout.total == 0) return B2DP_OK;
  DeviceGuard guard(r.device);
  StepArgs a;
  memset(&a, 0, sizeof a);
  for (int i = 0; i < n; ++i) a.out[i] = reinterpret_cast<float*>(r.base[i] + off);
  a.w = a.out[root];
  a.nvec = p->layout.total / 4;
  a.push = (r.rank == root) ? 1 : 0;
  fill_comm(p, r, &a.c);
  cudaStream_t s = (cudaStream_t)stream;
  auto go = [&](auto kernel) {
    const int grid = step_grid(kernel, r, a.nvec, 1, p->grid_cap);
    kernel<<<grid, kStepThreads, 0, s>>>(a, root);
    return cudaGetLastError();
  };
  cudaError_t e;
  switch (n) {
    case 2: e = go(broadcast_kernel<2>); break;
    case 3: e = go(broadcast_kernel<3>); break;
    case 4: e = go(broadcast_kernel<4>); break;
    case 5: e = go(broadcast_kernel<5>); break;
    case 6: e = go(broadcast_kernel<6>); break;
    case 7: e = go(broadcast_kernel<7>); break;
    default: e = go(broadcast_kernel<8>); break;
  }
  if (e != cudaSuccess) return fail(B2DP_ERR_CUDA, "broadcast launch: %s", cudaGetErrorString(e));
  return B2DP_OK;
}

extern "C" int b2dp_allreduce_small(b2dp_plan* p, int32_t dev, const float* in, float* out, int32_t count,
                                    float postscale, void* stream) {
  Range nvtx_range("b2dp::allreduce_small");
  RankState* rp = nullptr;
  if (int rc = get_rank(p, dev, &rp)) return rc;
  RankState& r = *rp;
  if (count < 0 || count > kSmallMax)
    return fail(count < 0 ? B2DP_ERR_INVALID : B2DP_ERR_TOO_LARGE, "allreduce_small: count %d outside 0..%d (use b2dp_allreduce for tensors)", count, kSmallMax);
  if (count == 0) return B2DP_OK;
  if (!in || !out) return fail(B2DP_ERR_INVALID, "allreduce_small: null pointer");
  if (!b2dp_plan_ready(p)) return fail(B2DP_ERR_STATE, "plan has unattached peers");
  if (int rc = check_poisoned(r)) return rc;
  const int n = p->ctx->nranks;
  DeviceGuard guard(r.device);
  cudaStream_t s = (cudaStream_t)stream;
  SmallArgs a;
  memset(&a, 0, sizeof a);
  for (int i = 0; i < n; ++i) a.slots[i] = reinterpret_cast<float*>(r.base[i] + p->map.small);
  a.in = in;
  a.out = out;
  a.count = count;
  a.half = (int)(r.small_calls++ & 1);
  a.postscale = postscale;
  fill_comm(p, r, &a.c);
  if (!p->ctx->emulated) {
    a.do_put = a.do_sum = 1;
    small_allreduce_kernel<<<1, 256, 0, s>>>(a);
    CUDA_TRY(cudaGetLastError());
    return B2DP_OK;
  }
  // emulated ranks (one GPU, no waiting kernels): deposit now; after the last rank has
  // deposited, run every rank's sum
  a.do_put = 1;
  a.do_sum = 0;
  small_allreduce_kernel<<<1, 256, 0, s>>>(a);
  CUDA_TRY(cudaGetLastError());
  r.small_out = out;
  r.small_count = count;
  r.small_post = postscale;
  if (r.rank == n - 1) {
    for (auto& q : p->rs) {
      SmallArgs b = a;
      b.do_put = 0;
      b.do_sum = 1;
      b.c.rank = q->rank;
      b.out = q->small_out;
      b.count = q->small_count;
      b.postscale = q->small_post;
      if (!b.out) return fail(B2DP_ERR_STATE, "allreduce_small: emulated rank %d never made its call", q->rank);
      small_allreduce_kernel<<<1, 256, 0, s>>>(b);
      CUDA_TRY(cudaGetLastError());
      q->small_out = nullptr;
    }
  }
  return B2DP_OK;
}

extern "C" float b2dp_adam_alpha(float lr, float beta1, float beta2, int64_t step) {
  // fp32 scalar arithmetic, one rounding per operation (Keras computes these as fp32 tensors)
  volatile float b1p = powf(beta1, (float)step);
  volatile float b2p = powf(beta2, (float)step);
  volatile float num = sqrtf(1.0f - b2p);
  volatile float den = 1.0f - b1p;
  volatile float a = lr * num;
  return a / den;
}
