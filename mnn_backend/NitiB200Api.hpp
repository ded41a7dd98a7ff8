// NitiB200Api.hpp — table of C-ABI entry points of libniti_b200.so (include/niti_b200.h), bound with
// dlopen/dlsym the way the reference binds its accelerator library: a struct of function pointers and an
// all-or-nothing `loaded` flag (execution-engine/source/backend/cpu/hexagon_implementation.cpp:52-113,
// include/MNN/hexagon/hexagon_nn/hexagon_implementation.h:24-130).
// The plug-in links no CUDA library: device memory, copies and kernels all go through this table.
#ifndef MNN_NITI_B200_API_HPP
#define MNN_NITI_B200_API_HPP
#include "niti_b200.h"

namespace MNN {
namespace B200 {

struct NitiB200Api {
    bool loaded = false;
    decltype(&niti_ctx_create)        ctx_create        = nullptr;
    decltype(&niti_ctx_destroy)       ctx_destroy       = nullptr;
    decltype(&niti_last_error)        last_error        = nullptr;
    decltype(&niti_conv_fwd_acc)      conv_fwd_acc      = nullptr;
    decltype(&niti_conv_fwd_fused)    conv_fwd_fused    = nullptr;
    decltype(&niti_conv_dx_acc)       conv_dx_acc       = nullptr;
    decltype(&niti_conv_dw_acc)       conv_dw_acc       = nullptr;
    decltype(&niti_matmul_nt_acc)     matmul_nt_acc     = nullptr;
    decltype(&niti_zero_u32)          zero_u32          = nullptr;
    decltype(&niti_absmax)            absmax            = nullptr;
    decltype(&niti_rescale_act)       rescale_act       = nullptr;
    decltype(&niti_rescale_grad)      rescale_grad      = nullptr;
    decltype(&niti_rescale_grad_split) rescale_grad_split = nullptr;
    decltype(&niti_rescale_dsp)       rescale_dsp       = nullptr;
    decltype(&niti_sgd_update)        sgd_update        = nullptr;
    decltype(&niti_relu)              relu              = nullptr;
    decltype(&niti_relu_grad)         relu_grad         = nullptr;
    decltype(&niti_maxpool)           maxpool           = nullptr;
    decltype(&niti_maxpool_grad)      maxpool_grad      = nullptr;
    decltype(&niti_loss_grad)         loss_grad         = nullptr;
    decltype(&niti_loss_value)        loss_value        = nullptr;
    decltype(&niti_onehot_labels)     onehot_labels     = nullptr;
    decltype(&niti_binary_i8)         binary_i8         = nullptr;
    decltype(&niti_zero_insert_c4)    zero_insert_c4    = nullptr;
    decltype(&niti_pad2d)             pad2d             = nullptr;
    decltype(&niti_gather_i8)         gather_i8         = nullptr;
    decltype(&niti_nc4hw4_to_nhwc)    nc4hw4_to_nhwc    = nullptr;
    decltype(&niti_nhwc_to_nc4hw4)    nhwc_to_nc4hw4    = nullptr;
    decltype(&niti_dev_alloc)         dev_alloc         = nullptr;
    decltype(&niti_dev_free)          dev_free          = nullptr;
    decltype(&niti_dev_memset)        dev_memset        = nullptr;
    decltype(&niti_copy_h2d)          copy_h2d          = nullptr;
    decltype(&niti_copy_d2h)          copy_d2h          = nullptr;
    decltype(&niti_copy_d2d)          copy_d2d          = nullptr;
    decltype(&niti_stream_sync)       stream_sync       = nullptr;
    decltype(&niti_launch_count)      launch_count      = nullptr;
};

// Loads libniti_b200.so once (path: $NITI_B200_LIB, else next to this plug-in, else the loader path).
// Returns a table with loaded == false when the library or any symbol is missing — there is no CPU
// fallback behind it: the RuntimeCreator then reports the backend as invalid.
const NitiB200Api& nitiApi();

} // namespace B200
} // namespace MNN
#endif
