// layout.cpp -- the gradient bucketing layer (host arithmetic only, no CUDA).
//
// Stands in for tf.distribute's _pack_tensors / group_by_size(bytes_per_pack)
// [reached from MirroredStrategy(), imagenet-resnet50-mirror.py:21, and from
// CommunicationOptions, imagenet-resnet50-multiworkers.py:20-21] and for Horovod's
// fusion buffer [hvd.DistributedOptimizer, imagenet-resnet50-hvd.py:101]: the K gradient
// tensors of one step are laid end to end into buckets, and every bucket is cut into
// nranks equal slices -- slice r is the part of the model rank r reduces, updates and
// republishes (the PS shard of imagenet-resnet50-ps.py:75-82).
#include <algorithm>
#include <cstring>

#include "internal.h"

namespace b2dp {

static thread_local std::string g_err;

int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

static inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

}  // namespace b2dp

using namespace b2dp;

extern "C" const char* b2dp_last_error(void) { return g_err.c_str(); }
extern "C" int b2dp_abi_version(void) { return 1; }

extern "C" int b2dp_layout_create(int32_t K, const int64_t* numel, int32_t nranks,
                                  int32_t wire_dtype, int64_t bucket_bytes, b2dp_layout** out) {
  if (!out) return fail(B2DP_ERR_INVALID, "layout_create: out is null");
  *out = nullptr;
  if (K < 0 || (K > 0 && !numel)) return fail(B2DP_ERR_INVALID, "layout_create: bad tensor list");
  if (nranks < 1 || nranks > kMaxRanks)
    return fail(B2DP_ERR_INVALID, "layout_create: nranks %d outside 1..%d", nranks, kMaxRanks);
  if (wire_dtype != B2DP_F32 && wire_dtype != B2DP_BF16)
    return fail(B2DP_ERR_INVALID, "layout_create: wire dtype %d is not f32/bf16", wire_dtype);
  if (bucket_bytes < 0) return fail(B2DP_ERR_INVALID, "layout_create: negative bucket_bytes");
  for (int32_t k = 0; k < K; ++k)
    if (numel[k] < 0 || numel[k] > (int64_t(1) << 40))
      return fail(B2DP_ERR_INVALID, "layout_create: tensor %d has numel %lld", k, (long long)numel[k]);

  auto* l = new b2dp_layout;
  l->K = K;
  l->nranks = nranks;
  l->wire = wire_dtype;
  l->bucket_bytes = bucket_bytes;
  l->numel.assign(numel, numel + K);
  l->offset.resize(K);
  l->bucket_of.resize(K);

  const int64_t slice_quant = kSliceAlign * nranks;
  int64_t cursor = 0;      // next free flat element
  int64_t shard = 0;
  Bucket cur;
  cur.offset = 0;
  int64_t cur_bytes = 0;   // fp32 bytes of the tensors in the open bucket
  bool open = false;

  auto close_bucket = [&]() {
    cur.elems = round_up(cursor - cur.offset, slice_quant);
    cur.slice = cur.elems / nranks;
    cur.shard_off = shard;
    shard += cur.slice;
    cursor = cur.offset + cur.elems;
    l->buckets.push_back(cur);
    open = false;
  };

  for (int32_t k = 0; k < K; ++k) {
    const int64_t bytes = 4 * numel[k];
    if (open && bucket_bytes > 0 && cur.count > 0 && cur_bytes + bytes > bucket_bytes) close_bucket();
    if (!open) {
      cur = Bucket();
      cur.first = k;
      cur.offset = cursor;
      cur_bytes = 0;
      open = true;
    }
    cursor = round_up(cursor, kTensorAlign);
    l->offset[k] = cursor;
    l->bucket_of[k] = (int32_t)l->buckets.size();
    cursor += numel[k];
    cur.count += 1;
    cur_bytes += bytes;
  }
  if (open) close_bucket();
  l->total = cursor;
  l->shard = shard;

  // Tile table: work is split by bytes, not by tensor -- 160 of ResNet-50's 214
  // tensors are <= 2048 elements while 10 hold 64% of the bytes.
  for (auto& b : l->buckets) {
    b.tile_first = (int64_t)l->tiles.size();
    for (int32_t k = b.first; k < b.first + b.count; ++k) {
      for (int64_t s = 0; s < numel[k]; s += kTileElems) {
        Tile t;
        t.tensor = k;
        t.src_off = s;
        t.dst_off = l->offset[k] + s;
        t.count = (uint32_t)std::min<int64_t>(kTileElems, numel[k] - s);
        l->tiles.push_back(t);
      }
    }
    b.tile_count = (int64_t)l->tiles.size() - b.tile_first;
  }
  *out = l;
  return B2DP_OK;
}

extern "C" void b2dp_layout_destroy(b2dp_layout* l) { delete l; }
extern "C" int32_t b2dp_layout_num_tensors(const b2dp_layout* l) { return l ? l->K : -1; }
extern "C" int32_t b2dp_layout_num_buckets(const b2dp_layout* l) { return l ? (int32_t)l->buckets.size() : -1; }
extern "C" int32_t b2dp_layout_nranks(const b2dp_layout* l) { return l ? l->nranks : -1; }
extern "C" int64_t b2dp_layout_total_elems(const b2dp_layout* l) { return l ? l->total : -1; }
extern "C" int64_t b2dp_layout_shard_elems(const b2dp_layout* l) { return l ? l->shard : -1; }

extern "C" int64_t b2dp_layout_tensor_offset(const b2dp_layout* l, int32_t k) {
  if (!l || k < 0 || k >= l->K) return fail(B2DP_ERR_INVALID, "tensor_offset: index %d out of range", k);
  return l->offset[k];
}

extern "C" int32_t b2dp_layout_tensor_bucket(const b2dp_layout* l, int32_t k) {
  if (!l || k < 0 || k >= l->K) return fail(B2DP_ERR_INVALID, "tensor_bucket: index %d out of range", k);
  return l->bucket_of[k];
}

extern "C" int b2dp_layout_bucket_info(const b2dp_layout* l, int32_t b, int64_t info[6]) {
  if (!l || !info || b < 0 || b >= (int32_t)l->buckets.size())
    return fail(B2DP_ERR_INVALID, "bucket_info: bucket %d out of range", b);
  const Bucket& k = l->buckets[b];
  info[0] = k.first;
  info[1] = k.count;
  info[2] = k.offset;
  info[3] = k.elems;
  info[4] = k.slice;
  info[5] = k.shard_off;
  return B2DP_OK;
}
