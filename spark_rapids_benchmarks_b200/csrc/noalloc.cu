// nds_join_probe_i64_noalloc: a hash-join lookup that performs no device or host allocation.
//
// Who needs this: the engine's normal path (nds_engine_execute) takes its hash tables from CUDA's
// stream-ordered pool, which is fine for a stand-alone process.  A host that embeds this library next to
// its own GPU memory manager cannot accept that: the spark-rapids plugin, for one, hands almost all of the
// device to its RMM pool when the executor starts, so a cudaMalloc issued from inside a library either fails
// with out-of-memory or takes memory the pool's spill accounting never sees.  The same holds for callers on a
// latency budget that must not touch the driver's allocator lock in the middle of a query.  For them the
// join's scratch (the open-addressing table) lives in a fixed array inside the loaded module.
//
// The price: a fixed ceiling (kNoAllocMaxBuild build keys), ~1.5 MB of device memory held by every GPU that
// loads the library until the process exits, and one table shared by all calls (see nds_b200.h).
#include <mutex>
#include "engine.h"
#include "../../include/nds_b200.h"

namespace nds {

namespace {

constexpr int kNoAllocMaxBuild = NDS_NOALLOC_MAX_BUILD_KEYS;      // 65536
constexpr uint32_t kNoAllocSlots = 2u * kNoAllocMaxBuild;          // load factor <= 0.5
constexpr size_t kKeysBytes = (size_t)kNoAllocSlots * sizeof(int64_t);
constexpr size_t kValsBytes = (size_t)kNoAllocSlots * sizeof(uint32_t);
constexpr size_t kFlagsOff = kKeysBytes + kValsBytes;
constexpr size_t kScratchBytes = kFlagsOff + 256;

// The scratch itself.  In host code this name is only a handle for cudaGetSymbolAddress - it is not a device
// address and must never be passed to a kernel directly.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_setup_mu;
unsigned char* g_scratch_addr[kMaxDevices] = {nullptr};

// Per device, with that device current: resolve the array's device address once.
unsigned char* scratch_for_current_device() {
  int dev = -1;
  NDS_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) fail("nds_join_probe_i64_noalloc: device index out of range");
  std::lock_guard<std::mutex> lock(g_setup_mu);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p)
      fail(std::string("nds_join_probe_i64_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
    g_scratch_addr[dev] = static_cast<unsigned char*>(p);
  }
  return g_scratch_addr[dev];
}

// Everything that must start clean for a call: every key slot empty, both flag words zero.
__global__ void noalloc_clear_kernel(int64_t* keys, uint32_t slots, int* flags) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < slots; i += gridDim.x * blockDim.x) keys[i] = kEmptyKey;
  if (blockIdx.x == 0 && threadIdx.x < 2) flags[threadIdx.x] = 0;
}

// Dense variant of join_probe_kernel (ops.cu): one output per probe key, -1 when nothing matches.
__global__ void noalloc_probe_dense_kernel(const int64_t* __restrict__ probe, int64_t n, HashTable ht, int32_t* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t k = probe[i];
    int32_t hit = -1;
    if (k != kEmptyKey) {
      uint32_t s = (uint32_t)mix64((uint64_t)k) & ht.mask;
      for (;;) {
        const int64_t cur = ht.keys[s];
        if (cur == k) { hit = (int32_t)ht.vals[s]; break; }
        if (cur == kEmptyKey) break;
        s = (s + 1) & ht.mask;
      }
    }
    out[i] = hit;
  }
}

}  // namespace

void join_probe_i64_noalloc(const int64_t* build_keys, int32_t n_build, const int64_t* probe_keys, int64_t n_probe,
                            int32_t* out_build_row, cudaStream_t stream) {
  if (n_build < 0 || n_probe < 0) fail("nds_join_probe_i64_noalloc: negative size");
  if (n_build > kNoAllocMaxBuild)
    fail("nds_join_probe_i64_noalloc: " + std::to_string(n_build) + " build keys exceed the fixed scratch table (maximum " +
         std::to_string(kNoAllocMaxBuild) + "); use nds_engine_execute, which sizes its table from the memory pool");
  if ((n_build > 0 && !build_keys) || (n_probe > 0 && (!probe_keys || !out_build_row)))
    fail("nds_join_probe_i64_noalloc: null pointer argument");
  unsigned char* base = scratch_for_current_device();
  HashTable ht;
  ht.keys = reinterpret_cast<int64_t*>(base);
  ht.vals = reinterpret_cast<uint32_t*>(base + kKeysBytes);
  ht.mask = kNoAllocSlots - 1;
  int* flags = reinterpret_cast<int*>(base + kFlagsOff);
  // the table still holds the previous call's keys: reset it on the caller's stream before building
  noalloc_clear_kernel<<<148, 256, 0, stream>>>(ht.keys, kNoAllocSlots, flags);
  NDS_CUDA(cudaGetLastError());
  if (n_build > 0) {
    ColRef key;
    key.data = build_keys;
    key.valid = nullptr;
    key.rowmap = nullptr;
    key.strbase = nullptr;
    key.dtype = DT_I64;
    key.pad = 0;
    launch_join_build(key, RowCount{nullptr, n_build}, ht, flags, stream);
  }
  if (n_probe > 0) {
    int64_t blocks = (n_probe + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    noalloc_probe_dense_kernel<<<(int)blocks, 256, 0, stream>>>(probe_keys, n_probe, ht, out_build_row);
    NDS_CUDA(cudaGetLastError());
  }
  int h_flags[2] = {0, 0};
  NDS_CUDA(cudaMemcpyAsync(h_flags, flags, sizeof h_flags, cudaMemcpyDeviceToHost, stream));
  NDS_CUDA(cudaStreamSynchronize(stream));
  if (h_flags[0] & 1) fail("nds_join_probe_i64_noalloc: duplicate keys on the build side (build keys must be unique)");
  if (h_flags[0] & 2) fail("nds_join_probe_i64_noalloc: a build key equals the reserved empty marker INT64_MIN");
}

}  // namespace nds

extern "C" NDS_API int nds_join_probe_i64_noalloc(const int64_t* build_keys_dev, int32_t n_build, const int64_t* probe_keys_dev,
                                                  int64_t n_probe, int32_t* out_build_row_dev, void* cuda_stream) {
  try {
    nds::join_probe_i64_noalloc(build_keys_dev, n_build, probe_keys_dev, n_probe, out_build_row_dev,
                                reinterpret_cast<cudaStream_t>(cuda_stream));
  } catch (const std::exception& ex) {
    nds::set_global_error(ex.what());
    cudaGetLastError();
    return 1;
  }
  return 0;
}
