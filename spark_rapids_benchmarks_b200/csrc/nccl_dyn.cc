#include "nccl_dyn.h"
#include <dlfcn.h>
#include <mutex>
#include <string>
#include "common.h"

namespace nds {

static NcclApi g_api;
static std::once_flag g_once;
static std::string g_load_error;

static void load() {
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    g_api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (g_api.lib) break;
  }
  if (!g_api.lib) { g_load_error = std::string("cannot load libnccl.so.2: ") + dlerror(); return; }
  auto sym = [&](const char* s) -> void* {
    void* p = dlsym(g_api.lib, s);
    if (!p && g_load_error.empty()) g_load_error = std::string("libnccl lacks ") + s;
    return p;
  };
  g_api.GetUniqueId = reinterpret_cast<int (*)(NcclUniqueId*)>(sym("ncclGetUniqueId"));
  g_api.CommInitRank = reinterpret_cast<int (*)(void**, int, NcclUniqueId, int)>(sym("ncclCommInitRank"));
  g_api.CommDestroy = reinterpret_cast<int (*)(void*)>(sym("ncclCommDestroy"));
  g_api.AllGather = reinterpret_cast<int (*)(const void*, void*, size_t, int, void*, cudaStream_t)>(sym("ncclAllGather"));
  g_api.Send = reinterpret_cast<int (*)(const void*, size_t, int, int, void*, cudaStream_t)>(sym("ncclSend"));
  g_api.Recv = reinterpret_cast<int (*)(void*, size_t, int, int, void*, cudaStream_t)>(sym("ncclRecv"));
  g_api.GroupStart = reinterpret_cast<int (*)()>(sym("ncclGroupStart"));
  g_api.GroupEnd = reinterpret_cast<int (*)()>(sym("ncclGroupEnd"));
  g_api.GetErrorString = reinterpret_cast<const char* (*)(int)>(sym("ncclGetErrorString"));
  g_api.GetVersion = reinterpret_cast<int (*)(int*)>(sym("ncclGetVersion"));
}

const NcclApi& nccl_api() {
  std::call_once(g_once, load);
  if (!g_load_error.empty()) fail("NCCL is not available: " + g_load_error);
  return g_api;
}

void nccl_check(int rc, const char* what) {
  if (rc == 0) return;
  const NcclApi& a = nccl_api();
  fail(std::string("NCCL error in ") + what + ": " + (a.GetErrorString ? a.GetErrorString(rc) : "?") + " (" + std::to_string(rc) + ")");
}

}  // namespace nds
