// Preload.cpp — optional LD_PRELOAD shim: run the UNMODIFIED reference binary
// (`runTrainDemo.out NITIInt8Train <mnist_dir>`) on the B200 backend with no source change.
//
// The demo hard-codes `exe->setGlobalExecutorConfig(MNN_FORWARD_CPU, config, 4)`
// (tools/train/source/demo/MnistUtils.cpp:41-43).  With
//     LD_PRELOAD=libMNN_niti_b200_preload.so MNN_NITI_B200_AS_DEFAULT=1 ./runTrainDemo.out NITIInt8Train dir
// the dynamic linker resolves that call to the definition below, which forwards to the real
// Executor::setGlobalExecutorConfig (next definition in search order, i.e. libMNN) with the forward
// type replaced by MNN_FORWARD_USER_0.  Everything above the backend (express, grad, optimizer, demo)
// runs as shipped.  Without the environment variable the call is forwarded untouched.
#include <dlfcn.h>
#include <cstdio>
#include <cstdlib>
#include <MNN/expr/Executor.hpp>
#include "B200Backend.hpp"

namespace MNN {
namespace Express {

void Executor::setGlobalExecutorConfig(MNNForwardType type, const BackendConfig& config, int numberThread) {
    typedef void (*RealFn)(Executor*, MNNForwardType, const BackendConfig&, int);
    static RealFn real = (RealFn)::dlsym(RTLD_NEXT, "_ZN3MNN7Express8Executor23setGlobalExecutorConfigE14MNNForwardTypeRKNS_13BackendConfigEi");
    if (nullptr == real) {
        ::fprintf(stderr, "[niti_b200] preload: cannot find the real Executor::setGlobalExecutorConfig\n");
        ::abort();
    }
    const char* env = ::getenv("MNN_NITI_B200_AS_DEFAULT");
    if (env && env[0] == '1' && type == MNN_FORWARD_CPU) {
        if (nullptr == MNNGetExtraRuntimeCreator(MNN_FORWARD_B200)) {
            ::fprintf(stderr, "[niti_b200] preload: B200 runtime unavailable (libniti_b200.so / sm_100 device missing); "
                              "refusing to fall back to the CPU\n");
            ::abort();
        }
        ::fprintf(stderr, "[niti_b200] preload: MNN_FORWARD_CPU -> MNN_FORWARD_USER_0 (B200 NITI backend)\n");
        type = MNN_FORWARD_B200;
    }
    real(this, type, config, numberThread);
}

} // namespace Express
} // namespace MNN
