#include "common.cuh"

#include <cstdlib>

namespace tfc {
std::string& last_error_ref() {
  static thread_local std::string s;
  return s;
}
std::atomic<int64_t> g_launches{0};
std::atomic<int> g_sm_budget{0};
std::atomic<int> g_coder_pack{0};
std::atomic<int> g_coder_lanes{0};
std::atomic<int> g_tiles_per_cta{0};
std::atomic<int> g_debug{0};
std::atomic<int> g_conv_cluster{0};
std::atomic<int> g_conv_stages{0};
// TFC_CONV_PAIR in the environment presets the CTA-pair mode (tests run the whole suite with it)
static int env_int(const char* name, int lo, int hi) {
  const char* e = getenv(name);
  if (!e || !*e) return 0;
  const int v = atoi(e);
  return (v < lo || v > hi) ? 0 : v;
}
std::atomic<int> g_conv_pair{env_int("TFC_CONV_PAIR", 0, 2)};
std::atomic<int> g_coder_table_kb{0};
std::atomic<int> g_coder_smem_kb{0};
}  // namespace tfc

extern "C" const char* tfc_last_error(void) { return tfc::last_error_ref().c_str(); }
extern "C" const char* tfc_version(void) { return "tfc_b200 0.1 (sm_100a)"; }
extern "C" int64_t tfc_launch_count(void) { return tfc::g_launches.load(); }

extern "C" int tfc_configure(int key, int value) {
  switch (key) {
    case TFC_CFG_SM_BUDGET:
      if (value < 0) return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "SM budget must be >= 0");
      tfc::g_sm_budget.store(value);
      return TFC_OK;
    case TFC_CFG_CODER_STREAMS_PER_CTA:
      // warps-per-stream kernels use min(value, 8); the lane-per-stream kernels value / 32 warps
      if (value < 0 || value > 512)
        return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "coder streams per CTA must be in [0, 512]");
      tfc::g_coder_pack.store(value);
      return TFC_OK;
    case TFC_CFG_TILES_PER_CTA:
      if (value < 0 || value > 4096)
        return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "tiles per CTA must be in [0, 4096]");
      tfc::g_tiles_per_cta.store(value);
      return TFC_OK;
    case TFC_CFG_CONV_CLUSTER:
      if (value != 0 && value != 1 && value != 2 && value != 4)
        return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "conv cluster size must be 0 (auto), 1, 2 or 4");
      tfc::g_conv_cluster.store(value);
      return TFC_OK;
    case TFC_CFG_CODER_LANES:
      if (value < 0 || value > 2)
        return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "coder lanes mode must be 0 (auto), 1 (off) or 2 (on)");
      tfc::g_coder_lanes.store(value);
      return TFC_OK;
    case TFC_CFG_CONV_STAGES:
      if (value != 0 && (value < 2 || value > 6))
        return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "conv stages must be 0 (auto) or 2..6");
      tfc::g_conv_stages.store(value);
      return TFC_OK;
    case TFC_CFG_CODER_TABLE_KB:
      if (value < 0 || value > 72)
        return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "coder table budget must be in [0, 72] KB");
      tfc::g_coder_table_kb.store(value);
      return TFC_OK;
    case TFC_CFG_CONV_PAIR:
      if (value < 0 || value > 2)
        return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "conv CTA-pair mode must be 0 (auto), 1 (off) or 2 (on)");
      tfc::g_conv_pair.store(value);
      return TFC_OK;
    case TFC_CFG_CODER_SMEM_KB:
      if (value != 0 && (value < 16 || value > 226))
        return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "coder shared-memory budget must be 0 or in [16, 226] KB");
      tfc::g_coder_smem_kb.store(value);
      return TFC_OK;
    case 99:   // kernel experiments (tools/conv_bench.py --debug): results are NOT valid when set
      tfc::g_debug.store(value);
      return TFC_OK;
    default:
      return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "unknown configuration key %d", key);
  }
}
