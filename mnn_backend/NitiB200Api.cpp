// NitiB200Api.cpp — dlopen/dlsym binding of libniti_b200.so (see NitiB200Api.hpp).
#include "NitiB200Api.hpp"
#include <dlfcn.h>
#include <cstdio>
#include <cstdlib>
#include <string>

namespace MNN {
namespace B200 {

static std::string dirOfThisLibrary() {
    Dl_info info;
    if (dladdr((const void*)&dirOfThisLibrary, &info) && info.dli_fname) {
        std::string p(info.dli_fname);
        auto pos = p.find_last_of('/');
        if (pos != std::string::npos) return p.substr(0, pos);
    }
    return ".";
}

static void* openLibrary() {
    if (const char* env = ::getenv("NITI_B200_LIB")) {
        if (void* h = ::dlopen(env, RTLD_NOW | RTLD_LOCAL)) return h;
        ::fprintf(stderr, "[niti_b200] cannot load $NITI_B200_LIB=%s: %s\n", env, ::dlerror());
        return nullptr;
    }
    const std::string here = dirOfThisLibrary();
    const std::string candidates[] = {
        here + "/libniti_b200.so",
        here + "/../mandheling_dsp_training_b200/libniti_b200.so",
        "libniti_b200.so",
    };
    for (auto& c : candidates) {
        if (void* h = ::dlopen(c.c_str(), RTLD_NOW | RTLD_LOCAL)) return h;
    }
    ::fprintf(stderr, "[niti_b200] libniti_b200.so not found (set NITI_B200_LIB): %s\n", ::dlerror());
    return nullptr;
}

const NitiB200Api& nitiApi() {
    static NitiB200Api api;
    static bool tried = false;
    if (tried) return api;
    tried = true;
    void* h = openLibrary();
    if (!h) return api;
    bool ok = true;
#define NITI_BIND(field, sym)                                                   \
    api.field = reinterpret_cast<decltype(api.field)>(::dlsym(h, #sym));        \
    if (!api.field) { ::fprintf(stderr, "[niti_b200] missing symbol %s\n", #sym); ok = false; }
    NITI_BIND(ctx_create, niti_ctx_create)
    NITI_BIND(ctx_destroy, niti_ctx_destroy)
    NITI_BIND(last_error, niti_last_error)
    NITI_BIND(conv_fwd_acc, niti_conv_fwd_acc)
    NITI_BIND(conv_fwd_fused, niti_conv_fwd_fused)
    NITI_BIND(conv_dx_acc, niti_conv_dx_acc)
    NITI_BIND(conv_dw_acc, niti_conv_dw_acc)
    NITI_BIND(matmul_nt_acc, niti_matmul_nt_acc)
    NITI_BIND(zero_u32, niti_zero_u32)
    NITI_BIND(absmax, niti_absmax)
    NITI_BIND(rescale_act, niti_rescale_act)
    NITI_BIND(rescale_grad, niti_rescale_grad)
    NITI_BIND(rescale_grad_split, niti_rescale_grad_split)
    NITI_BIND(rescale_dsp, niti_rescale_dsp)
    NITI_BIND(sgd_update, niti_sgd_update)
    NITI_BIND(relu, niti_relu)
    NITI_BIND(relu_grad, niti_relu_grad)
    NITI_BIND(maxpool, niti_maxpool)
    NITI_BIND(maxpool_grad, niti_maxpool_grad)
    NITI_BIND(loss_grad, niti_loss_grad)
    NITI_BIND(loss_value, niti_loss_value)
    NITI_BIND(onehot_labels, niti_onehot_labels)
    NITI_BIND(binary_i8, niti_binary_i8)
    NITI_BIND(zero_insert_c4, niti_zero_insert_c4)
    NITI_BIND(pad2d, niti_pad2d)
    NITI_BIND(gather_i8, niti_gather_i8)
    NITI_BIND(nc4hw4_to_nhwc, niti_nc4hw4_to_nhwc)
    NITI_BIND(nhwc_to_nc4hw4, niti_nhwc_to_nc4hw4)
    NITI_BIND(dev_alloc, niti_dev_alloc)
    NITI_BIND(dev_free, niti_dev_free)
    NITI_BIND(dev_memset, niti_dev_memset)
    NITI_BIND(copy_h2d, niti_copy_h2d)
    NITI_BIND(copy_d2h, niti_copy_d2h)
    NITI_BIND(copy_d2d, niti_copy_d2d)
    NITI_BIND(stream_sync, niti_stream_sync)
    NITI_BIND(launch_count, niti_launch_count)
#undef NITI_BIND
    api.loaded = ok;            // all-or-nothing, like hexagon_implementation.cpp:104 `interface_loaded`
    return api;
}

} // namespace B200
} // namespace MNN
