// librnb_cupti.so: kernel timeline through the CUPTI Activity API (declared in include/rnb_cupti.h).
#include <cupti.h>
#include <cxxabi.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <string>
#include <vector>

#include "../../include/rnb_cupti.h"

namespace {

struct KernelRecord {
  std::string name;
  uint64_t start, end;
  uint32_t device, stream;
  int32_t grid[3], block[3];
};

std::mutex g_mu;
std::vector<KernelRecord> g_records;
std::string g_error;
bool g_enabled = false;
bool g_callbacks = false;
size_t g_dropped = 0;

constexpr size_t kBufferBytes = 4u << 20;  // one activity buffer; CUPTI asks for more as needed
constexpr size_t kBufferAlign = 8;

int fail(const char* what, CUptiResult r) {
  const char* s = nullptr;
  cuptiGetResultString(r, &s);
  std::lock_guard<std::mutex> lk(g_mu);
  g_error = std::string(what) + ": " + (s ? s : "unknown CUPTI error");
  return -1;
}

std::string demangle(const char* name) {
  if (!name) return "";
  int status = 0;
  char* d = abi::__cxa_demangle(name, nullptr, nullptr, &status);
  std::string out = (status == 0 && d) ? d : name;
  free(d);
  return out;
}

void CUPTIAPI buffer_requested(uint8_t** buffer, size_t* size, size_t* max_records) {
  void* p = nullptr;
  if (posix_memalign(&p, kBufferAlign, kBufferBytes) != 0) p = nullptr;
  *buffer = static_cast<uint8_t*>(p);
  *size = p ? kBufferBytes : 0;
  *max_records = 0;  // as many as fit
}

void CUPTIAPI buffer_completed(CUcontext ctx, uint32_t stream_id, uint8_t* buffer, size_t /*size*/,
                               size_t valid) {
  if (valid > 0) {
    std::vector<KernelRecord> batch;
    CUpti_Activity* rec = nullptr;
    while (cuptiActivityGetNextRecord(buffer, valid, &rec) == CUPTI_SUCCESS) {
      if (rec->kind != CUPTI_ACTIVITY_KIND_CONCURRENT_KERNEL && rec->kind != CUPTI_ACTIVITY_KIND_KERNEL)
        continue;
      const CUpti_ActivityKernel9* k = reinterpret_cast<const CUpti_ActivityKernel9*>(rec);
      KernelRecord r;
      r.name = demangle(k->name);
      r.start = k->start;
      r.end = k->end;
      r.device = k->deviceId;
      r.stream = k->streamId;
      r.grid[0] = k->gridX; r.grid[1] = k->gridY; r.grid[2] = k->gridZ;
      r.block[0] = k->blockX; r.block[1] = k->blockY; r.block[2] = k->blockZ;
      batch.push_back(std::move(r));
    }
    size_t dropped = 0;
    if (cuptiActivityGetNumDroppedRecords(ctx, stream_id, &dropped) != CUPTI_SUCCESS) dropped = 0;
    std::lock_guard<std::mutex> lk(g_mu);
    g_dropped += dropped;
    for (auto& r : batch) g_records.push_back(std::move(r));
  }
  free(buffer);
}

}  // namespace

extern "C" {

const char* rnb_cupti_last_error(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  static thread_local std::string copy;
  copy = g_error;
  return copy.c_str();
}

int rnb_cupti_initialize(void) {
  CUptiResult r;
  if (!g_callbacks) {
    r = cuptiActivityRegisterCallbacks(buffer_requested, buffer_completed);
    if (r != CUPTI_SUCCESS) return fail("cuptiActivityRegisterCallbacks", r);
    g_callbacks = true;
  }
  if (!g_enabled) {
    r = cuptiActivityEnable(CUPTI_ACTIVITY_KIND_CONCURRENT_KERNEL);
    if (r != CUPTI_SUCCESS) return fail("cuptiActivityEnable(CONCURRENT_KERNEL)", r);
    g_enabled = true;
  }
  return 0;
}

int rnb_cupti_flush(void) {
  const CUptiResult r = cuptiActivityFlushAll(0);
  if (r != CUPTI_SUCCESS) return fail("cuptiActivityFlushAll", r);
  return 0;
}

int rnb_cupti_count(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  return (int)g_records.size();
}

int rnb_cupti_get(int index, char* name, int name_cap, uint64_t* start_ns, uint64_t* end_ns,
                  uint32_t* device, uint32_t* stream, int32_t grid3[3], int32_t block3[3]) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (index < 0 || index >= (int)g_records.size()) {
    g_error = "rnb_cupti_get: index out of range";
    return -2;
  }
  const KernelRecord& r = g_records[index];
  if (name && name_cap > 0) snprintf(name, (size_t)name_cap, "%s", r.name.c_str());
  if (start_ns) *start_ns = r.start;
  if (end_ns) *end_ns = r.end;
  if (device) *device = r.device;
  if (stream) *stream = r.stream;
  if (grid3) memcpy(grid3, r.grid, sizeof r.grid);
  if (block3) memcpy(block3, r.block, sizeof r.block);
  return 0;
}

int rnb_cupti_clear(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  g_records.clear();
  g_dropped = 0;
  return 0;
}

int rnb_cupti_finalize(void) {
  if (g_enabled) {
    (void)cuptiActivityFlushAll(0);
    const CUptiResult r = cuptiActivityDisable(CUPTI_ACTIVITY_KIND_CONCURRENT_KERNEL);
    if (r != CUPTI_SUCCESS) return fail("cuptiActivityDisable", r);
    g_enabled = false;
  }
  return 0;
}

}  // extern "C"
