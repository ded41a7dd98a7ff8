// B200NitiElementwise.cpp — the int8 element-wise ops that sit between two contractions of the step
// (SURVEY.md §8 f1), so activations stay in HBM instead of bouncing through WrapExecution copies:
//   OpType_NITI_Relu_Int8      NITI_CPURelu_Int8::onExecute      source/backend/cpu/NITI_CPURelu_Int8.cpp:29-58
//   OpType_NITI_ReluGrad_Int8  NITI_CPUReluGrad_Int8::onExecute  NITI_CPUReluGrad_Int8.cpp:28-62  {x, dy} -> dx = x > 0 ? dy : 0
//   OpType_NITI_Maxpool_Int8   NITI_Maxpool_Int8::onExecute      NITI_Maxpool_Int8.cpp:24-70,127-181 {x NHWC, exp} -> {y NHWC, exp}
// Relu / ReluGrad are layout-agnostic (the reference walks the flat buffer), pooling is NHWC on both sides.
#include "B200Backend.hpp"
#include "core/TensorUtils.hpp"
#include "MNN_generated.h"

namespace MNN {
namespace B200 {
namespace {

inline int8_t* dev(const Tensor* t) { return (int8_t*)(uintptr_t)t->deviceId(); }
inline bool isInt8(const Tensor* t) { return t->getType().code == halide_type_int && t->getType().bits == 8; }

class B200NitiRelu : public Execution {
public:
    B200NitiRelu(Backend* bn, bool grad) : Execution(bn), mGrad(grad) {}
    virtual ErrorCode onExecute(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto bn = static_cast<B200Backend*>(backend());
        const size_t n = (size_t)outputs[0]->size();         // whole buffer, padded NC4HW4 lanes included
        int rc = mGrad ? bn->api().relu_grad(bn->ctx(), dev(inputs[0]), dev(inputs[1]), dev(outputs[0]), n, bn->stream())
                       : bn->api().relu(bn->ctx(), dev(inputs[0]), dev(outputs[0]), n, bn->stream());
        return toErrorCode(rc, bn, mGrad ? "niti_relu_grad" : "niti_relu");
    }
private:
    bool mGrad;
};

class B200NitiMaxpool : public Execution {
public:
    B200NitiMaxpool(Backend* bn, const NITI_Pool_Int8* pool) : Execution(bn), mPool(pool) {}
    virtual ErrorCode onResize(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto x = inputs[0], y = outputs[0];
        if (TensorUtils::getDescribe(x)->dimensionFormat != MNN_DATA_FORMAT_NHWC) return NOT_SUPPORT;
        if (mPool->isGlobal() || mPool->padX() != 0 || mPool->padY() != 0 || mPool->padType() == PoolPadType_SAME) return NOT_SUPPORT;
        if (mPool->kernelX() != mPool->kernelY() || mPool->strideX() != mPool->strideY()) return NOT_SUPPORT;
        mK = mPool->kernelX(); mS = mPool->strideX();
        if (mK > x->height() || mK > x->width()) return NOT_SUPPORT;
        if ((x->height() - mK) / mS + 1 != y->height() || (x->width() - mK) / mS + 1 != y->width()) return NOT_SUPPORT;
        return NO_ERROR;
    }
    virtual ErrorCode onExecute(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto bn = static_cast<B200Backend*>(backend());
        auto x = inputs[0];
        int rc = bn->api().maxpool(bn->ctx(), dev(x), x->batch(), x->height(), x->width(), x->channel(), mK, mS,
                                   dev(outputs[0]), bn->stream());
        if (rc == NITI_OK && outputs.size() > 1 && inputs.size() > 1) {      // exponent passes through (:168)
            rc = bn->api().copy_d2d(bn->ctx(), dev(outputs[1]), dev(inputs[1]), 1, bn->stream());
        }
        return toErrorCode(rc, bn, "niti_maxpool");
    }
private:
    const NITI_Pool_Int8* mPool;
    int mK = 2, mS = 2;
};

class ReluCreator : public B200Backend::Creator {
public:
    virtual Execution* onCreate(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs,
                                const MNN::Op* op, Backend* backend) const override {
        const bool grad = op->type() == OpType_NITI_ReluGrad_Int8;
        if (inputs.size() < (grad ? 2u : 1u) || !isInt8(inputs[0])) return nullptr;
        return new B200NitiRelu(backend, grad);
    }
};
class MaxpoolCreator : public B200Backend::Creator {
public:
    virtual Execution* onCreate(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs,
                                const MNN::Op* op, Backend* backend) const override {
        auto pool = op->main_as_NITI_Pool_Int8();
        if (nullptr == pool || inputs.empty() || !isInt8(inputs[0])) return nullptr;
        return new B200NitiMaxpool(backend, pool);
    }
};

typedef ReluCreator ReluGradCreator;
REGISTER_B200_OP_CREATOR(ReluCreator, OpType_NITI_Relu_Int8);
REGISTER_B200_OP_CREATOR(ReluGradCreator, OpType_NITI_ReluGrad_Int8);
REGISTER_B200_OP_CREATOR(MaxpoolCreator, OpType_NITI_Maxpool_Int8);

} // namespace
} // namespace B200
} // namespace MNN
