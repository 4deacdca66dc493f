// Optional per-kernel timing with CUDA events on the launching stream (bench.py's roofline leg).
// Off by default; when on, every launcher brackets its kernel with a pair of events.
#include <vector>

#include "common.cuh"

struct ProfSlot {
  const char* name;
  std::vector<cudaEvent_t> ev;   // start, stop, start, stop, ...
};
static ProfSlot g_slots[32];
static int g_nslots = 0;
static bool g_prof_on = false;

bool uuo_prof_enabled() { return g_prof_on; }

static ProfSlot* slot_for(const char* name) {
  for (int i = 0; i < g_nslots; ++i)
    if (g_slots[i].name == name || strcmp(g_slots[i].name, name) == 0) return &g_slots[i];
  if (g_nslots >= 32) return nullptr;
  g_slots[g_nslots].name = name;
  return &g_slots[g_nslots++];
}

void uuo_prof_mark(const char* name, cudaStream_t st) {
  if (!g_prof_on) return;
  ProfSlot* s = slot_for(name);
  if (!s) return;
  cudaEvent_t e;
  if (cudaEventCreate(&e) != cudaSuccess) return;
  cudaEventRecord(e, st);
  s->ev.push_back(e);
}

extern "C" int uuo_profile_enable(int on) {
  for (int i = 0; i < g_nslots; ++i) {
    for (cudaEvent_t e : g_slots[i].ev) cudaEventDestroy(e);
    g_slots[i].ev.clear();
  }
  g_nslots = 0;
  g_prof_on = on != 0;
  return UUO_OK;
}

extern "C" int uuo_profile_read(int slot, char* name, int name_cap, float* total_ms, int* count) {
  if (slot < 0 || slot >= g_nslots) return g_nslots;
  ProfSlot& s = g_slots[slot];
  float tot = 0.f;
  int n = 0;
  for (size_t i = 0; i + 1 < s.ev.size(); i += 2) {
    float ms = 0.f;
    UUO_CUDA(cudaEventSynchronize(s.ev[i + 1]));
    UUO_CUDA(cudaEventElapsedTime(&ms, s.ev[i], s.ev[i + 1]));
    tot += ms;
    ++n;
  }
  if (name && name_cap > 0) { strncpy(name, s.name, name_cap - 1); name[name_cap - 1] = 0; }
  if (total_ms) *total_ms = tot;
  if (count) *count = n;
  return g_nslots;
}
