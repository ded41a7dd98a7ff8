// B200NitiElementwise.cpp — the int8 element-wise ops that sit between two contractions of the step
// (SURVEY.md §8 f1), so activations stay in HBM instead of bouncing through WrapExecution copies:
//   OpType_NITI_Relu_Int8      NITI_CPURelu_Int8::onExecute      source/backend/cpu/NITI_CPURelu_Int8.cpp:29-58
//   OpType_NITI_ReluGrad_Int8  NITI_CPUReluGrad_Int8::onExecute  NITI_CPUReluGrad_Int8.cpp:28-62  {x, dy} -> dx = x > 0 ? dy : 0
//   OpType_NITI_Maxpool_Int8   NITI_Maxpool_Int8::onExecute      NITI_Maxpool_Int8.cpp:24-70,127-181 {x NHWC, exp} -> {y NHWC, exp}
//   OpType_NITI_PoolGrad_Int8  NITI_CPUMaxPoolGrad_Int8::onExecute NITI_CPUPoolGrad_Int8.cpp:21-74  {x, y, dy} NC4HW4 -> dx
//   OpType_NITI_LeftPoolGrad_Int8 NITI_CPULeftPoolGrad_Int8::onExecute NITI_CPULeftPoolGrad_Int8.cpp:19-52 zero insertion, NC4HW4
//   OpType_NITI_PAD_Int8       NITI_Pad_Int8::onExecute           NITI_Pad_Int8.cpp:26-66  NCHW planes
//   OpType_NITI_LOSS_Grad_Int8 NITI_CPULossGrad_Int8::onExecute   NITI_CPULossGrad_Int8.cpp:82-201 {logits, ascale, one-hot, dLoss} -> dLogits
//   OpType_BinaryOp (int8 x int8) NITI_CPUBinary_Int8::onExecute  CPUBinary.cpp:246-330: ADD / SUB / MUL clipped to +-127
// Relu / ReluGrad are layout-agnostic (the reference walks the flat buffer), pooling is NHWC on both sides.
#include "B200Backend.hpp"
#include "core/TensorUtils.hpp"
#include <memory>
#include "core/Macro.h"
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
    B200NitiMaxpool(Backend* bn, const NITI_Pool_Int8* pool) : Execution(bn), mPool(pool), mK(pool->kernelX()), mS(pool->strideX()) {}
    // An unsupported configuration must be refused by the CREATOR (nullptr -> the executor's CPU backup backend,
    // express/Executor.cpp:521-529); an error from onResize aborts the whole compute (:570-573).
    static bool supported(const NITI_Pool_Int8* pool, const Tensor* x, const Tensor* y) {
        if (TensorUtils::getDescribe(x)->dimensionFormat != MNN_DATA_FORMAT_NHWC) return false;
        if (pool->isGlobal() || pool->padX() != 0 || pool->padY() != 0 || pool->padType() == PoolPadType_SAME) return false;
        if (pool->kernelX() != pool->kernelY() || pool->strideX() != pool->strideY() || pool->strideX() <= 0) return false;
        const int k = pool->kernelX(), s = pool->strideX();
        if (k > x->height() || k > x->width()) return false;
        return (x->height() - k) / s + 1 == y->height() && (x->width() - k) / s + 1 == y->width();
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

// max-pool gradient on [slab][h][w][4] blocks: NC4HW4 with slab = batch * ceil(c/4) (each slab is an independent 4-channel
// image, which is how the reference walks it), or NHWC with slab = batch
class B200NitiPoolGrad : public Execution {
public:
    B200NitiPoolGrad(Backend* bn, const NITI_Pool_Int8* pool) : Execution(bn), mK(pool->kernelX()), mS(pool->strideX()) {}
    // the reference walks batch * ceil(c/4) slabs of [h][w][4] whatever the tensors' declared format is (:30-38); with
    // c % 4 == 0 every format has exactly that many bytes, so the same walk is reproduced
    static bool slabView(const Tensor* t) { return TensorUtils::getDescribe(t)->dimensionFormat == MNN_DATA_FORMAT_NC4HW4 || t->channel() % 4 == 0; }
    static bool supported(const NITI_Pool_Int8* pool, const Tensor* x, const Tensor* y, const Tensor* dy, const Tensor* dx) {
        if (!slabView(x) || !slabView(y) || !slabView(dy) || !slabView(dx)) return false;
        if (TensorUtils::getDescribe(x)->dimensionFormat == MNN_DATA_FORMAT_NHWC) return false;      // width()/height() would mean other axes
        if (pool->isGlobal() || pool->padX() != 0 || pool->padY() != 0 || pool->padType() == PoolPadType_SAME) return false;
        if (pool->kernelX() != pool->kernelY() || pool->strideX() != pool->strideY() || pool->strideX() <= 0) return false;
        const int k = pool->kernelX(), s = pool->strideX();
        if (k > x->height() || k > x->width()) return false;
        return (x->height() - k) / s + 1 == dy->height() && (x->width() - k) / s + 1 == dy->width();
    }
    virtual ErrorCode onExecute(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto bn = static_cast<B200Backend*>(backend());
        auto x = inputs[0];
        int rc = bn->api().maxpool_grad(bn->ctx(), dev(x), dev(inputs[1]), dev(inputs[2]), x->batch() * UP_DIV(x->channel(), 4),
                                        x->height(), x->width(), 4, mK, mS, dev(outputs[0]), bn->stream());
        return toErrorCode(rc, bn, "niti_maxpool_grad");
    }
private:
    int mK, mS;
};

class B200NitiLeftPoolGrad : public Execution {
public:
    B200NitiLeftPoolGrad(Backend* bn, const NITI_Pool_Int8* pool) : Execution(bn), mSx(pool->strideX()), mSy(pool->strideY()) {}
    static bool supported(const NITI_Pool_Int8* pool, const Tensor* in, const Tensor* out) {
        if (!B200NitiPoolGrad::slabView(in) || !B200NitiPoolGrad::slabView(out)) return false;       // same slab walk, :28-36
        if (TensorUtils::getDescribe(in)->dimensionFormat == MNN_DATA_FORMAT_NHWC) return false;
        const int sx = pool->strideX(), sy = pool->strideY();
        return sx > 0 && sy > 0 && (out->height() - 1) / sy < in->height() && (out->width() - 1) / sx < in->width();
    }
    virtual ErrorCode onExecute(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto bn = static_cast<B200Backend*>(backend());
        auto in = inputs[0], out = outputs[0];
        int rc = bn->api().zero_insert_c4(bn->ctx(), dev(in), in->batch() * UP_DIV(in->channel(), 4), in->height(), in->width(),
                                          out->height(), out->width(), mSy, mSx, dev(out), bn->stream());
        return toErrorCode(rc, bn, "niti_zero_insert_c4");
    }
private:
    int mSx, mSy;
};

class B200NitiPad : public Execution {
public:
    B200NitiPad(Backend* bn, int pad) : Execution(bn), mPad(pad) {}
    // the reference walks stride(0) / stride(1) planes of [h][w]: plain NCHW on both sides
    static bool supported(int pad, const Tensor* in, const Tensor* out) {
        if (TensorUtils::getDescribe(in)->dimensionFormat != MNN_DATA_FORMAT_NCHW ||
            TensorUtils::getDescribe(out)->dimensionFormat != MNN_DATA_FORMAT_NCHW) return false;
        if (in->dimensions() != 4 || out->dimensions() != 4 || in->batch() != out->batch() || in->channel() != out->channel()) return false;
        return pad >= 0 && out->height() - 2 * pad <= in->height() && out->height() - 2 * pad >= 0 && out->width() >= in->width() + pad;
    }
    virtual ErrorCode onExecute(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto bn = static_cast<B200Backend*>(backend());
        auto in = inputs[0], out = outputs[0];
        int rc = bn->api().pad2d(bn->ctx(), dev(in), in->batch() * in->channel(), in->height(), in->width(), mPad, out->height(),
                                 out->width(), dev(out), bn->stream());
        return toErrorCode(rc, bn, "niti_pad2d");
    }
private:
    int mPad;
};

// integer softmax-cross-entropy gradient: the one-hot int32 target becomes class indices on the device, then niti_loss_grad
class B200NitiLossGrad : public Execution {
public:
    explicit B200NitiLossGrad(Backend* bn) : Execution(bn) {}
    static bool supported(const Tensor* logits, const Tensor* target) {
        const int b = logits->batch();
        if (b <= 0 || logits->elementSize() <= 0 || logits->elementSize() % b != 0) return false;
        if (TensorUtils::getDescribe(logits)->dimensionFormat == MNN_DATA_FORMAT_NC4HW4) return false;   // [b][classes] rows expected
        return target->getType().bits == 32 && target->elementSize() == logits->elementSize();
    }
    virtual ErrorCode onResize(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto logits = inputs[0];
        mBatch = logits->batch();
        mClasses = logits->elementSize() / mBatch;
        mLabels.reset(Tensor::createDevice<int32_t>({mBatch}));
        if (!backend()->onAcquireBuffer(mLabels.get(), Backend::DYNAMIC)) return OUT_OF_MEMORY;
        backend()->onReleaseBuffer(mLabels.get(), Backend::DYNAMIC);
        return NO_ERROR;
    }
    virtual ErrorCode onExecute(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto bn = static_cast<B200Backend*>(backend());
        int32_t* labels = (int32_t*)(uintptr_t)mLabels->deviceId();
        int rc = bn->api().onehot_labels(bn->ctx(), (const int32_t*)(uintptr_t)inputs[2]->deviceId(), mBatch, mClasses, labels, bn->stream());
        if (rc == NITI_OK)
            rc = bn->api().loss_grad(bn->ctx(), dev(inputs[0]), dev(inputs[1]), labels, mBatch, mClasses, dev(outputs[0]), bn->stream());
        return toErrorCode(rc, bn, "niti_loss_grad");
    }
private:
    std::shared_ptr<Tensor> mLabels;
    int mBatch = 0, mClasses = 0;
};

// the float loss scalar (NITI_CPULoss_Int8.cpp:68-131): same operands as the gradient op, one float out
class B200NitiLoss : public Execution {
public:
    explicit B200NitiLoss(Backend* bn) : Execution(bn) {}
    virtual ErrorCode onResize(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto logits = inputs[0];
        mBatch = logits->batch();
        mClasses = logits->elementSize() / mBatch;
        mTargetClasses = inputs[2]->elementSize() / mBatch;      // one-hot rows may be narrower than the logit rows (:98-108)
        mLabels.reset(Tensor::createDevice<int32_t>({mBatch}));
        mRows.reset(Tensor::createDevice<float>({mBatch}));
        if (!backend()->onAcquireBuffer(mLabels.get(), Backend::DYNAMIC) || !backend()->onAcquireBuffer(mRows.get(), Backend::DYNAMIC)) return OUT_OF_MEMORY;
        backend()->onReleaseBuffer(mLabels.get(), Backend::DYNAMIC);
        backend()->onReleaseBuffer(mRows.get(), Backend::DYNAMIC);
        return NO_ERROR;
    }
    virtual ErrorCode onExecute(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto bn = static_cast<B200Backend*>(backend());
        int32_t* labels = (int32_t*)(uintptr_t)mLabels->deviceId();
        int rc = bn->api().onehot_labels(bn->ctx(), (const int32_t*)(uintptr_t)inputs[2]->deviceId(), mBatch, mTargetClasses, labels, bn->stream());
        if (rc == NITI_OK)
            rc = bn->api().loss_value(bn->ctx(), dev(inputs[0]), dev(inputs[1]), labels, mBatch, mClasses, mClasses,
                                      (float*)(uintptr_t)mRows->deviceId(), (float*)(uintptr_t)outputs[0]->deviceId(), bn->stream());
        return toErrorCode(rc, bn, "niti_loss_value");
    }
private:
    std::shared_ptr<Tensor> mLabels, mRows;
    int mBatch = 0, mClasses = 0, mTargetClasses = 0;
};

class B200NitiBinary : public Execution {
public:
    B200NitiBinary(Backend* bn, int op) : Execution(bn), mOp(op) {}
    virtual ErrorCode onExecute(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto bn = static_cast<B200Backend*>(backend());
        // element counts as CPUBackend::getTensorSize counts them (NC4HW4 channel lanes rounded up), :283-297
        const size_t n0 = lanes(inputs[0]), n1 = lanes(inputs[1]);
        const int bc = n0 == n1 ? -1 : (n0 == 1 ? 0 : 1);
        const size_t n = bc == 0 ? n1 : n0;
        int rc = bn->api().binary_i8(bn->ctx(), mOp, dev(inputs[0]), dev(inputs[1]), n, bc, dev(outputs[0]), bn->stream());
        return toErrorCode(rc, bn, "niti_binary_i8");
    }
private:
    static size_t lanes(const Tensor* t) {
        size_t n = 1;
        const bool c4 = TensorUtils::getDescribe(t)->dimensionFormat == MNN_DATA_FORMAT_NC4HW4;
        for (int i = 0; i < t->dimensions(); ++i) n *= (size_t)((c4 && i == 1) ? UP_DIV(t->length(i), 4) * 4 : t->length(i));
        return n;
    }
    int mOp;
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
        if (nullptr == pool || inputs.empty() || outputs.empty() || !isInt8(inputs[0])) return nullptr;
        if (!B200NitiMaxpool::supported(pool, inputs[0], outputs[0])) return nullptr;
        return new B200NitiMaxpool(backend, pool);
    }
};

class PoolGradCreator : public B200Backend::Creator {
public:
    virtual Execution* onCreate(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs,
                                const MNN::Op* op, Backend* backend) const override {
        auto pool = op->main_as_NITI_Pool_Int8();
        if (nullptr == pool || inputs.empty() || !isInt8(inputs[0])) return nullptr;
        if (outputs.empty()) return nullptr;
        if (op->type() == OpType_NITI_LeftPoolGrad_Int8)
            return B200NitiLeftPoolGrad::supported(pool, inputs[0], outputs[0]) ? new B200NitiLeftPoolGrad(backend, pool) : nullptr;
        if (inputs.size() < 3 || !B200NitiPoolGrad::supported(pool, inputs[0], inputs[1], inputs[2], outputs[0])) return nullptr;
        return new B200NitiPoolGrad(backend, pool);
    }
};
class PadCreator : public B200Backend::Creator {
public:
    virtual Execution* onCreate(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs,
                                const MNN::Op* op, Backend* backend) const override {
        auto pad = op->main_as_NITI_PAD_Int8();
        if (nullptr == pad || inputs.empty() || outputs.empty() || !isInt8(inputs[0])) return nullptr;
        if (!B200NitiPad::supported(pad->pad(), inputs[0], outputs[0])) return nullptr;
        return new B200NitiPad(backend, pad->pad());
    }
};
class LossGradCreator : public B200Backend::Creator {
public:
    virtual Execution* onCreate(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs,
                                const MNN::Op* op, Backend* backend) const override {
        if (inputs.size() < 3 || outputs.empty() || !isInt8(inputs[0]) || !isInt8(inputs[1])) return nullptr;
        if (!B200NitiLossGrad::supported(inputs[0], inputs[2])) return nullptr;
        return new B200NitiLossGrad(backend);
    }
};
class LossCreator : public B200Backend::Creator {
public:
    virtual Execution* onCreate(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs,
                                const MNN::Op* op, Backend* backend) const override {
        if (inputs.size() < 3 || outputs.empty() || !isInt8(inputs[0]) || !isInt8(inputs[1])) return nullptr;
        auto logits = inputs[0]; auto target = inputs[2];
        const int b = logits->batch();
        if (b <= 0 || logits->elementSize() <= 0 || logits->elementSize() % b != 0 || logits->elementSize() / b < 2) return nullptr;
        if (TensorUtils::getDescribe(logits)->dimensionFormat == MNN_DATA_FORMAT_NC4HW4) return nullptr;      // [b][classes] rows expected
        if (target->getType().bits != 32 || target->elementSize() % b != 0 || target->elementSize() > logits->elementSize()) return nullptr;
        return new B200NitiLoss(backend);
    }
};
class BinaryCreator : public B200Backend::Creator {
public:
    virtual Execution* onCreate(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs,
                                const MNN::Op* op, Backend* backend) const override {
        auto b = op->main_as_BinaryOp();
        if (nullptr == b || inputs.size() < 2 || !isInt8(inputs[0]) || !isInt8(inputs[1])) return nullptr;   // other types: CPU backup
        const int type = b->opType();
        const int code = type == BinaryOpOperation_ADD ? 0 : (type == BinaryOpOperation_SUB ? 1 : (type == BinaryOpOperation_MUL ? 2 : -1));
        if (code < 0) return nullptr;
        return new B200NitiBinary(backend, code);
    }
};

typedef ReluCreator ReluGradCreator;
typedef PoolGradCreator LeftPoolGradCreator;
REGISTER_B200_OP_CREATOR(PoolGradCreator, OpType_NITI_PoolGrad_Int8);
REGISTER_B200_OP_CREATOR(LeftPoolGradCreator, OpType_NITI_LeftPoolGrad_Int8);
REGISTER_B200_OP_CREATOR(PadCreator, OpType_NITI_PAD_Int8);
REGISTER_B200_OP_CREATOR(LossGradCreator, OpType_NITI_LOSS_Grad_Int8);
REGISTER_B200_OP_CREATOR(LossCreator, OpType_NITI_LOSS_Int8);
REGISTER_B200_OP_CREATOR(BinaryCreator, OpType_BinaryOp);
REGISTER_B200_OP_CREATOR(ReluCreator, OpType_NITI_Relu_Int8);
REGISTER_B200_OP_CREATOR(ReluGradCreator, OpType_NITI_ReluGrad_Int8);
REGISTER_B200_OP_CREATOR(MaxpoolCreator, OpType_NITI_Maxpool_Int8);

} // namespace
} // namespace B200
} // namespace MNN
