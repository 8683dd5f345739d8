// extern "C" surface of libndsb200.so (declared in include/nds_b200.h).
#include <mutex>
#include "engine.h"
#include <memory>
#include <unordered_map>
#include "../../include/nds_b200.h"

namespace nds {
static std::mutex g_mu;
static std::string g_error;
void set_global_error(const std::string& s) {
  std::lock_guard<std::mutex> l(g_mu);
  g_error = s;
}
}  // namespace nds

struct nds_engine {
  nds::Engine* impl;
  // parsed plans by their JSON text: a statement that is run again (every step of a power run) skips the parse, like a
  // prepared statement.  Bounded; cleared when full.
  std::unordered_map<std::string, std::shared_ptr<const nds::Json>> plans;
};
static constexpr size_t kPlanCacheEntries = 256;

#define NDS_API_BEGIN try {
#define NDS_API_END(e)                                                  \
  }                                                                     \
  catch (const std::exception& ex) {                                    \
    if (e) (e)->impl->last_error = ex.what();                           \
    else nds::set_global_error(ex.what());                              \
    cudaGetLastError();                                                 \
    return 1;                                                           \
  }                                                                     \
  catch (...) {                                                         \
    if (e) (e)->impl->last_error = "unknown internal error";            \
    else nds::set_global_error("unknown internal error");               \
    cudaGetLastError();                                                 \
    return 1;                                                           \
  }                                                                     \
  return 0;

static int copy_out(const std::string& s, char* buf, int64_t buflen) {
  if (!buf || buflen <= 0) return 2;
  if ((int64_t)s.size() + 1 > buflen) return 3;
  memcpy(buf, s.c_str(), s.size() + 1);
  return 0;
}

extern "C" {

NDS_API const char* nds_version(void) { return "nds_b200 0.1 (sm_100a)"; }

NDS_API const char* nds_last_error(void) {
  static thread_local std::string copy;
  std::lock_guard<std::mutex> l(nds::g_mu);
  copy = nds::g_error;
  return copy.c_str();
}

NDS_API int nds_engine_create(const char* cfg_json, nds_engine** out) {
  nds_engine* none = nullptr;
  NDS_API_BEGIN
  if (!out) nds::fail("nds_engine_create: out is null");
  nds::Json cfg = nds::JsonParser(cfg_json && *cfg_json ? cfg_json : "{}").parse();
  nds::Engine* impl = new nds::Engine(cfg);
  *out = new nds_engine{impl};
  NDS_API_END(none)
}

NDS_API void nds_engine_destroy(nds_engine* e) {
  if (!e) return;
  delete e->impl;
  delete e;
}

NDS_API const char* nds_engine_last_error(nds_engine* e) { return e ? e->impl->last_error.c_str() : nds_last_error(); }

NDS_API int nds_engine_register_table(nds_engine* e, const char* name, const char* path, const char* format) {
  if (!e) return 2;
  NDS_API_BEGIN
  e->impl->register_table(name, path, format ? format : "parquet");
  NDS_API_END(e)
}

NDS_API int nds_engine_register_arrow(nds_engine* e, const char* name, struct ArrowSchema* schema, struct ArrowArray* array) {
  if (!e) return 2;
  NDS_API_BEGIN
  e->impl->register_arrow(name, schema, array);
  NDS_API_END(e)
}

NDS_API int nds_engine_drop_table(nds_engine* e, const char* name) {
  if (!e) return 2;
  NDS_API_BEGIN
  e->impl->drop_table(name);
  NDS_API_END(e)
}

NDS_API int nds_engine_table_info(nds_engine* e, const char* name, char* buf, int64_t buflen) {
  if (!e) return 2;
  NDS_API_BEGIN
  int rc = copy_out(e->impl->table_info_json(name), buf, buflen);
  if (rc) nds::fail("nds_engine_table_info: buffer too small");
  NDS_API_END(e)
}

NDS_API int nds_engine_execute(nds_engine* e, const char* plan_json, struct ArrowSchema* out_schema, struct ArrowArray* out_array) {
  if (!e) return 2;
  NDS_API_BEGIN
  if (!plan_json || !out_schema || !out_array) nds::fail("nds_engine_execute: null argument");
  std::shared_ptr<const nds::Json> plan;
  std::string text(plan_json);
  auto it = e->plans.find(text);
  if (it != e->plans.end()) {
    plan = it->second;
  } else {
    plan = std::make_shared<const nds::Json>(nds::JsonParser(plan_json).parse());
    if (e->plans.size() >= kPlanCacheEntries) e->plans.clear();
    e->plans.emplace(std::move(text), plan);
  }
  e->impl->execute(*plan, out_schema, out_array);
  NDS_API_END(e)
}

NDS_API int nds_engine_comm_unique_id(nds_engine* e, char* out, int64_t out_len) {
  if (!e) return 2;
  NDS_API_BEGIN
  if (!out || out_len < NDS_COMM_ID_BYTES) nds::fail("nds_engine_comm_unique_id: buffer of NDS_COMM_ID_BYTES needed");
  std::string id = e->impl->comm_unique_id();
  memcpy(out, id.data(), NDS_COMM_ID_BYTES);
  NDS_API_END(e)
}

NDS_API int nds_engine_comm_init(nds_engine* e, const char* id, int64_t id_len, int rank, int world) {
  if (!e) return 2;
  NDS_API_BEGIN
  if (!id || id_len != NDS_COMM_ID_BYTES) nds::fail("nds_engine_comm_init: id must be the NDS_COMM_ID_BYTES bytes rank 0 obtained");
  e->impl->comm_init(std::string(id, (size_t)id_len), rank, world);
  NDS_API_END(e)
}

NDS_API int nds_engine_last_stats(nds_engine* e, char* buf, int64_t buflen) {
  if (!e) return 2;
  NDS_API_BEGIN
  int rc = copy_out(e->impl->stats_json(), buf, buflen);
  if (rc) nds::fail("nds_engine_last_stats: buffer too small");
  NDS_API_END(e)
}

}  // extern "C"
