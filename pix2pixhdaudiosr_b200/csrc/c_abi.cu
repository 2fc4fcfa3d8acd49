// Error slot + version for libp2p_b200.so (include/p2p_b200.h).
#include "common.cuh"
#include <mutex>
#include <set>
#include <utility>

namespace p2p {
char* error_slot() {
    static thread_local char buf[512] = {0};
    return buf;
}
long long& launch_counter() {
    static long long n = 0;
    return n;
}
void note_kernel_for_carveout(const void* fn) {
    static std::mutex mu;
    static std::set<std::pair<const void*, int>> seen;
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lock(mu);
    if (seen.insert({fn, dev}).second)
        cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
}
int& deterministic_flag() {
    static int f = -1;
    return f;
}
}  // namespace p2p

extern "C" const char* p2p_last_error(void) { return p2p::error_slot(); }
extern "C" int p2p_abi_version(void) { return 3; }
extern "C" int p2p_set_deterministic(int on) {      // negative only queries; returns the old value
    const int old = p2p::deterministic() ? 1 : 0;
    if (on >= 0) p2p::deterministic_flag() = on ? 1 : 0;
    return old;
}
extern "C" long long p2p_kernel_launches(void) { return p2p::launch_counter(); }
