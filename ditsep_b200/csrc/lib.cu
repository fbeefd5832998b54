// Library-level state of libditsep_b200.so: last-error text, ABI version, launch counter.
#include <mutex>

#include "common.cuh"

namespace ditsep {

static thread_local char g_err[1024] = "";
std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

}  // namespace ditsep

extern "C" int ditsep_abi_version(void) { return 3; }  // 2: ragged-batch lengths on the pc_* entries, box_muller; 3: ditsep_sde_params.elem_scale, pc_langevin, energy_envelope
extern "C" const char* ditsep_last_error(void) { return ditsep::g_err; }
extern "C" int64_t ditsep_launch_count(void) { return ditsep::g_launches.load(); }
extern "C" void ditsep_note_launches(int64_t n) { ditsep::g_launches.fetch_add(n); }
