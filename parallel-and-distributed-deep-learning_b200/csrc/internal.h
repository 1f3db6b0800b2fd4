// internal.h -- shared between the host-side pieces of libb200dp.so (not installed).
#pragma once

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/b200dp.h"

namespace b2dp {

// ---- error plumbing ---------------------------------------------------------------
int fail(int code, const char* fmt, ...);  // records the thread-local message, returns code

// ---- layout -----------------------------------------------------------------------
constexpr int64_t kTensorAlign = 32;   // elements; every tensor starts on a 128-byte (f32) line
constexpr int64_t kSliceAlign = 256;   // elements; every rank slice is a multiple of this
constexpr int32_t kTileElems = 8192;   // pack/unpack work unit (one block)
constexpr int kMaxRanks = 8;           // one NVSwitch box

struct Tile {          // one unit of pack/unpack work: a run inside ONE tensor
  int64_t src_off;     // element offset inside the tensor
  int64_t dst_off;     // element offset inside the flat buffer
  int32_t tensor;
  uint32_t count;      // <= kTileElems
};

struct Bucket {
  int32_t first = 0, count = 0;      // tensor range
  int64_t offset = 0;                // flat element offset of the bucket
  int64_t elems = 0;                 // padded length, = nranks * slice
  int64_t slice = 0;                 // per-rank elements
  int64_t shard_off = 0;             // offset of this bucket inside a rank's state shard
  int64_t tile_first = 0, tile_count = 0;
};

}  // namespace b2dp

struct b2dp_layout {
  int32_t K = 0, nranks = 1, wire = B2DP_F32;
  int64_t bucket_bytes = 0;
  std::vector<int64_t> numel, offset;
  std::vector<int32_t> bucket_of;
  std::vector<b2dp::Bucket> buckets;
  std::vector<b2dp::Tile> tiles;
  int64_t total = 0;  // padded flat elements
  int64_t shard = 0;  // state elements per rank
};

namespace b2dp {

// ---- VMM (driver API through cudaGetDriverEntryPoint; no link-time libcuda) --------
struct VmmBlock {
  unsigned long long handle = 0;  // CUmemGenericAllocationHandle
  unsigned long long va = 0;      // CUdeviceptr where it is mapped in this process
  size_t size = 0;
  int owner_device = -1;          // device holding the physical memory (-1: imported peer)
  bool mapped = false;
};

int vmm_init();  // resolves the driver entry points once; B2DP_OK or error
int vmm_granularity(int device, size_t* gran);
// Allocates `size` (rounded up by the caller) physical bytes on `device`, maps them and
// grants read/write to every device in access_devices.
int vmm_alloc(int device, size_t size, const int* access_devices, int n_access, VmmBlock* out);
int vmm_export_fd(const VmmBlock& b, int* fd);
// Imports a peer's allocation from an fd and maps it for `device`.
int vmm_import_fd(int fd, size_t size, int device, VmmBlock* out);
void vmm_free(VmmBlock* b);

// NVSwitch multicast objects (Path B); every function returns B2DP_OK or an error
int mc_supported(int device);
int mc_granularity(int ndev, size_t size, size_t* gran);
int mc_create(int ndev, size_t size, unsigned long long* handle);
int mc_export_fd(unsigned long long handle, int* fd);
int mc_import_fd(int fd, unsigned long long* handle);
int mc_add_device(unsigned long long handle, int device);
int mc_bind(unsigned long long handle, const VmmBlock& mem, size_t size);
int mc_map(unsigned long long handle, size_t size, const int* devs, int ndev, unsigned long long* va_out);
void mc_release(unsigned long long handle, unsigned long long va, size_t size);

}  // namespace b2dp
