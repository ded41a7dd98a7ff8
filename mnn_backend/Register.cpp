// Register.cpp — registers the B200 runtime creator with MNN when the plug-in is loaded.
//
// Same mechanism as the reference's CUDA backend (source/backend/cuda/Register.cpp:12-35): a static
// initializer calls MNNInsertExtraRuntimeCreator (source/core/Backend.hpp:289-298) — here for the free
// plug-in slot MNN_FORWARD_USER_0.  A process selects it with
//     Executor::getGlobalExecutor()->setGlobalExecutorConfig(MNN_FORWARD_USER_0, config, threads);
// after dlopen("libMNN_niti_b200.so"), or links the plug-in.  See INTEGRATION.md.
#include "B200Backend.hpp"

namespace MNN {
namespace B200 {

static bool registerB200Runtime() {
    static B200RuntimeCreator* creator = new B200RuntimeCreator;
    // needCheck = true: MNNGetExtraRuntimeCreator then calls onValid() and hides the backend when
    // libniti_b200.so is missing (source/core/Backend.cpp:28-46) — the executor prints its own
    // "Error to find creator" message instead of silently computing elsewhere.
    return MNNInsertExtraRuntimeCreator(MNN_FORWARD_B200, creator, true);
}

static const bool gRegistered = registerB200Runtime();

} // namespace B200
} // namespace MNN

extern "C" {
// exported for hosts that prefer an explicit call over a static initializer (and for the symbol test)
int MNNNitiB200Registered() { return MNN::B200::gRegistered ? 1 : 0; }
int MNNNitiB200ForwardType() { return (int)MNN_FORWARD_B200; }
// CUDA kernels launched so far by the live B200 runtime (tests use it to prove the ops ran on the device)
long long MNNNitiB200LaunchCount() { return MNN::B200::launchCountOfLastRuntime(); }
// Executions the B200 backend created for this op type so far (0: every op of that type went to the CPU backup backend)
// the same for op types that went to the CPU backup backend; opType < 0: writes up to `cap` {type, count} pairs into `out`
// and returns how many types there are
long long MNNNitiB200DeclinedOps(int opType, long long* out, int cap) {
    auto& m = MNN::B200::declinedCounts();
    if (opType >= 0) { auto it = m.find(opType); return it == m.end() ? 0 : it->second; }
    int i = 0;
    for (auto& kv : m) { if (out && i < cap) { out[2 * i] = kv.first; out[2 * i + 1] = kv.second; } ++i; }
    return i;
}
long long MNNNitiB200CreatedOps(int opType) {
    auto& m = MNN::B200::createdCounts();
    auto it = m.find(opType);
    return it == m.end() ? 0 : it->second;
}
}
