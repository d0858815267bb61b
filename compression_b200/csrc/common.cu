#include "common.cuh"

namespace tfc {
std::string& last_error_ref() {
  static thread_local std::string s;
  return s;
}
std::atomic<int64_t> g_launches{0};
}  // namespace tfc

extern "C" const char* tfc_last_error(void) { return tfc::last_error_ref().c_str(); }
extern "C" const char* tfc_version(void) { return "tfc_b200 0.1 (sm_100a)"; }
extern "C" int64_t tfc_launch_count(void) { return tfc::g_launches.load(); }
