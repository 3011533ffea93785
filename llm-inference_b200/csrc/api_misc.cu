// Library identity + launch accounting.
#include <atomic>
#include "common.cuh"
#include "launch_count.h"

namespace b200 {
static std::atomic<uint64_t> g_launches{0};
void count_launch(uint64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace b200

extern "C" int b200_abi_version(void) { return 1; }
extern "C" const char* b200_build_info(void) {
    return "b200_llama sm_100a (compute_100a) built " __DATE__ " " __TIME__;
}
extern "C" uint64_t b200_launch_count(void) {
    return b200::g_launches.load(std::memory_order_relaxed);
}
extern "C" void b200_launch_count_reset(void) { b200::g_launches.store(0); }
