// B200NitiExecution.cpp — Executions + creators for the NITI int8 contraction op types.
//
// Each Execution takes the operands exactly as the reference's graph hands them to its CPU
// executions (same input order, shapes and layouts) and produces the same outputs:
//   OpType_NITI_CONV_Int8          NITI_Conv_Int8::onExecute          source/backend/cpu/NITI_Conv_Int8.cpp:162-310
//   OpType_NITI_DeCONV_Int8        NITI_DeConv_Int8::onExecute        NITI_DeConv_Int8.cpp:187-332
//   OpType_NITI_GradientCONV_Int8  NITI_GradientConv_Int8::onExecute  NITI_GradientConv_Int8.cpp:165-298
//   OpType_NITI_GradientSplitCONV_Int8 NITI_GradientSplitConv_Int8::onExecute NITI_GradientSplitConv_Int8.cpp:195-386
//   OpType_NITI_MatMul_Int8        NITI_Matmul_Int8::onExecute        NITI_Matmul_Int8.cpp:140-231
// The transposed / rotated / packed operand forms produced by the grad rule
// (tools/train/source/grad/NITI_Conv_Int8_Grad.cpp:27-191) are undone by one strided gather per
// operand (niti_gather_i8) so the kernels see their natural NHWC / OHWI operands; phase 1
// (contraction -> int32 + max|acc|) and phase 2 (shared exponent + PSTO) are the C-ABI calls.
#include <memory>
#include <vector>
#include "B200Backend.hpp"
#include "core/TensorUtils.hpp"
#include "core/Macro.h"
#include "MNN_generated.h"

namespace MNN {
namespace B200 {

namespace {

inline MNN_DATA_FORMAT formatOf(const Tensor* t) { return TensorUtils::getDescribe(t)->dimensionFormat; }
inline int8_t* dev(const Tensor* t) { return (int8_t*)(uintptr_t)t->deviceId(); }
inline bool isInt8(const Tensor* t) { return t->getType().code == halide_type_int && t->getType().bits == 8; }

// scratch device memory taken from the backend at onResize and handed back for reuse before it
// returns, the reference's own pattern (NITI_Conv_Int8.cpp:118-157)
class Scratch {
public:
    explicit Scratch(Backend* bn) : mBackend(bn) {}
    void* get(size_t bytes) {
        std::shared_ptr<Tensor> t(Tensor::createDevice<int8_t>({(int)((bytes + 255) / 256 * 256)}));
        if (!mBackend->onAcquireBuffer(t.get(), Backend::DYNAMIC)) { mFailed = true; return nullptr; }
        mTensors.emplace_back(t);
        return (void*)(uintptr_t)t->deviceId();
    }
    void releaseAll() {
        for (auto& t : mTensors) mBackend->onReleaseBuffer(t.get(), Backend::DYNAMIC);
    }
    void clear() { mTensors.clear(); mFailed = false; }
    bool failed() const { return mFailed; }
private:
    Backend* mBackend;
    std::vector<std::shared_ptr<Tensor>> mTensors;
    bool mFailed = false;
};

struct Gather {                       // one niti_gather_i8 call, prepared at onResize
    const int8_t* src = nullptr; int8_t* dst = nullptr;
    int64_t base = 0; int32_t dims[5] = {1, 1, 1, 1, 1}; int64_t strides[5] = {0, 0, 0, 0, 0};
    int splitHi = -1, splitLo = -1, splitLimit = 0;
    void set(int k, int dim, int64_t stride) { dims[k] = dim; strides[k] = stride; }
    int run(const B200Backend* bn) const {
        return bn->api().gather_i8(bn->ctx(), src, base, dst, dims, strides, splitHi, splitLo, splitLimit, bn->stream());
    }
};

// activation tensor [n, c, h, w] (any MNN format) -> contiguous NHWC [n][h*w][c]
Gather gatherToNHWC(const Tensor* t, int n, int c, int hw, int8_t* dst) {
    Gather g; g.src = dev(t); g.dst = dst;
    switch (formatOf(t)) {
        case MNN_DATA_FORMAT_NC4HW4: {                       // [c/4][n][hw][4]; c % 4 == 0 here
            g.set(0, n, (int64_t)hw * 4); g.set(1, hw, 4); g.set(2, c / 4, (int64_t)n * hw * 4); g.set(3, 4, 1);
            break;
        }
        case MNN_DATA_FORMAT_NHWC: g.set(0, n, (int64_t)hw * c); g.set(1, hw, c); g.set(2, c, 1); break;
        default:                   g.set(0, n, (int64_t)c * hw); g.set(1, hw, 1); g.set(2, c, hw); break;   // NCHW
    }
    return g;
}

// contiguous NHWC [n][h*w][c] -> activation tensor [n, c, h, w] in its MNN format (NC4HW4 lanes >= c are zeroed)
Gather gatherFromNHWC(const int8_t* src, const Tensor* t, int n, int c, int hw) {
    Gather g; g.src = src; g.dst = dev(t);
    switch (formatOf(t)) {
        case MNN_DATA_FORMAT_NC4HW4: {
            const int c4 = UP_DIV(c, 4);
            g.set(0, c4, 4); g.set(1, n, (int64_t)hw * c); g.set(2, hw, c); g.set(3, 4, 1);
            g.splitHi = 0; g.splitLo = 3; g.splitLimit = c;
            break;
        }
        case MNN_DATA_FORMAT_NHWC: g.set(0, n, (int64_t)hw * c); g.set(1, hw, c); g.set(2, c, 1); break;
        default:                   g.set(0, n, (int64_t)hw * c); g.set(1, c, 1); g.set(2, hw, c); break;    // NCHW
    }
    return g;
}

struct ConvGeometry {
    int stride = 1, dil = 1, pad = 0, kh = 1, kw = 1;
    bool ok = false;
};

// symmetric stride / dilation / padding is the C ABI's contract (niti_conv_desc); anything else is
// reported as unsupported so the executor can place the op on its CPU backup backend.
ConvGeometry readGeometry(const Convolution2DCommon* common, const Tensor* input, const Tensor* output) {
    ConvGeometry g;
    if (common->group() != 1) return g;
    if (common->strideX() != common->strideY() || common->dilateX() != common->dilateY()) return g;
    g.stride = common->strideX(); g.dil = common->dilateX();
    g.kh = common->kernelY(); g.kw = common->kernelX();
    int padX = common->padX(), padY = common->padY();        // ConvolutionCommon::convolutionPad (core/ConvolutionCommon.cpp:550-569)
    if (common->padMode() == PadMode_SAME) {
        padX = ((output->width() - 1) * g.stride + (g.kw - 1) * g.dil + 1 - input->width()) / 2;
        padY = ((output->height() - 1) * g.stride + (g.kh - 1) * g.dil + 1 - input->height()) / 2;
    } else if (nullptr != common->pads() && common->pads()->size() >= 2) {
        padX = common->pads()->data()[1]; padY = common->pads()->data()[0];
    }
    if (padX != padY || padX < 0) return g;
    g.pad = padX;
    const int oh = (input->height() + 2 * g.pad - ((g.kh - 1) * g.dil + 1)) / g.stride + 1;
    const int ow = (input->width() + 2 * g.pad - ((g.kw - 1) * g.dil + 1)) / g.stride + 1;
    g.ok = oh == output->height() && ow == output->width();
    return g;
}

class NitiExecution : public Execution {
public:
    explicit NitiExecution(Backend* bn) : Execution(bn), mScratch(bn) {}
protected:
    B200Backend* bn() const { return static_cast<B200Backend*>(backend()); }
    ErrorCode check(int rc, const char* what) const { return toErrorCode(rc, bn(), what); }
    Scratch mScratch;
    std::vector<Gather> mPre, mPost;
    ErrorCode runAll(const std::vector<Gather>& gs, const char* what) const {
        for (auto& g : gs) {
            int rc = g.run(bn());
            if (rc != NITI_OK) return check(rc, what);
        }
        return NO_ERROR;
    }
};

// ------------------------------------------------------------------------------------------------
// OpType_NITI_CONV_Int8: {x, w [oc,c,kh,kw] NCHW, ascale, wscale} -> {y, exponent}
// OpType_NITI_DeCONV_Int8: {dy (pre-padded), wT [c,oc,kh,kw] NCHW, wscale} -> {dx}: the reference runs it as a
//   forward convolution of dy with rot180(wT) (NITI_DeConv_Int8.cpp:179-219) and so do we.
class B200NitiConv : public NitiExecution {
public:
    B200NitiConv(Backend* bn, const Convolution2DCommon* common, bool deconv)
        : NitiExecution(bn), mCommon(common), mDeconv(deconv) {}

    virtual ErrorCode onResize(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto x = inputs[0], w = inputs[1], y = outputs[0];
        mPre.clear(); mPost.clear(); mScratch.clear();
        if (!isInt8(x) || !isInt8(w) || formatOf(w) != MNN_DATA_FORMAT_NCHW) return NOT_SUPPORT;
        auto geo = readGeometry(mCommon, x, y);
        if (!geo.ok) return NOT_SUPPORT;
        const int n = x->batch(), c = x->channel(), h = x->height(), wd = x->width();
        const int oc = y->channel(), oh = y->height(), ow = y->width();
        if (w->length(0) != oc || w->length(1) != c || w->length(2) != geo.kh || w->length(3) != geo.kw) return NOT_SUPPORT;
        if (formatOf(x) == MNN_DATA_FORMAT_NC4HW4 && c % 4 != 0 && c != 1) return NOT_SUPPORT;   // quirk Q6
        mDesc = niti_conv_desc{n, c, h, wd, oc, geo.kh, geo.kw, geo.stride, geo.pad, geo.dil};
        mOutCount = (size_t)n * oh * ow * oc;
        // exponent-prediction slot of the fused kernel: the express graph (and with it every Execution) is rebuilt each step, so
        // the slot is derived from the op's geometry plus its rank among the ops of that geometry in this backend's graph — the
        // same layer lands on the same slot in every step, and two layers of equal geometry do not share a predictor; a
        // collision only costs prediction accuracy, never results (include/niti_b200.h niti_set_speculate)
        {
            unsigned long long hsh = 1469598103934665603ull;
            for (int v : {n, c, h, wd, oc, geo.kh, geo.kw, geo.stride, geo.pad, geo.dil, mDeconv ? 1 : 0}) { hsh ^= (unsigned long long)(unsigned)v; hsh *= 1099511628211ull; }
            if (mOccurrence < 0 || hsh != mGeoKey) { mGeoKey = hsh; mOccurrence = bn()->occurrenceOf(hsh); }    // once per Execution and geometry: a second onResize keeps it
            hsh ^= (unsigned long long)(unsigned)mOccurrence; hsh *= 1099511628211ull;
            mSlot = 1 + (int)(hsh % 255ull);
        }

        const int kk = geo.kh * geo.kw;
        mX = (int8_t*)mScratch.get((size_t)n * h * wd * c);
        mW = (int8_t*)mScratch.get((size_t)oc * kk * c);
        mAcc = (int32_t*)mScratch.get(mOutCount * 4);
        mY = (int8_t*)mScratch.get(mOutCount);
        mMax = (uint32_t*)mScratch.get(256);
        if (mScratch.failed()) return OUT_OF_MEMORY;

        if (formatOf(x) == MNN_DATA_FORMAT_NC4HW4 && c % 4 != 0) {
            mRaggedIn = true;                                 // C == 1 stem: lane 0 of the single channel block
        } else {
            mRaggedIn = false;
            mPre.emplace_back(gatherToNHWC(x, n, c, h * wd, mX));
        }
        Gather gw; gw.src = dev(w); gw.dst = mW;              // OIHW -> OHWI (rot180 for DeCONV)
        gw.set(0, oc, (int64_t)c * kk); gw.set(3, c, kk);
        if (mDeconv) { gw.set(1, geo.kh, -(int64_t)geo.kw); gw.set(2, geo.kw, -1); gw.base = (int64_t)kk - 1; }
        else         { gw.set(1, geo.kh, geo.kw);           gw.set(2, geo.kw, 1); }
        mPre.emplace_back(gw);
        mPost.emplace_back(gatherFromNHWC(mY, y, n, oc, oh * ow));
        mScratch.releaseAll();
        return NO_ERROR;
    }

    virtual ErrorCode onExecute(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto& a = bn()->api(); auto ctx = bn()->ctx(); auto st = bn()->stream();
        if (mRaggedIn) {
            int rc = a.nc4hw4_to_nhwc(ctx, dev(inputs[0]), mDesc.n, mDesc.c, mDesc.h, mDesc.w, mX, st);
            if (rc != NITI_OK) return check(rc, "nc4hw4_to_nhwc");
        }
        auto code = runAll(mPre, "operand gather");
        if (code != NO_ERROR) return code;
        // contraction + max|acc| + shared exponent + PSTO behind ONE call: a single launch wherever the output tiles fit in
        // tensor memory (csrc/conv_chain.cu), the two-phase route through mAcc otherwise — same bytes either way
        int rc = a.zero_u32(ctx, mMax, 1, st);
        if (rc == NITI_OK) {
            if (mDeconv) rc = a.conv_fwd_fused(ctx, &mDesc, mX, mW, nullptr, nullptr, mY, nullptr, 0, mAcc, mMax, mSlot, st);
            else rc = a.conv_fwd_fused(ctx, &mDesc, mX, mW, dev(inputs[2]), dev(inputs[3]), mY, dev(outputs[1]), 0, mAcc, mMax, mSlot, st);
        }
        if (rc != NITI_OK) return check(rc, "niti_conv_fwd_fused");
        return runAll(mPost, "result gather");
    }

private:
    const Convolution2DCommon* mCommon;
    bool mDeconv;
    bool mRaggedIn = false;
    int mSlot = 0;
    int mOccurrence = -1;
    unsigned long long mGeoKey = 0;
    niti_conv_desc mDesc;
    size_t mOutCount = 0;
    int8_t *mX = nullptr, *mW = nullptr, *mY = nullptr;
    int32_t* mAcc = nullptr;
    uint32_t* mMax = nullptr;
};

// ------------------------------------------------------------------------------------------------
// OpType_NITI_GradientCONV_Int8 / GradientSplitCONV: {xT [C,N,H,W], dyT [OC,N,OH,OW], ascale, wscale} -> {dw [C,OC,KH,KW]}
// The reference convolves xT with dyT as the kernel (kernelX/Y = OH/OW, inputCount = N); that is the
// weight gradient of the stride-1 convolution x * w -> dy, which is what niti_conv_dw_acc computes from the
// natural operands.  Stride-2 layers arrive here already zero-inserted (NITI_LeftPoolGrad), i.e. as stride 1.
class B200NitiGradientConv : public NitiExecution {
public:
    B200NitiGradientConv(Backend* bn, const Convolution2DCommon* common, bool split)
        : NitiExecution(bn), mCommon(common), mSplit(split) {}

    virtual ErrorCode onResize(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto xT = inputs[0], dyT = inputs[1], dw = outputs[0];
        mPre.clear(); mPost.clear(); mScratch.clear();
        if (!isInt8(xT) || !isInt8(dyT)) return NOT_SUPPORT;
        auto geo = readGeometry(mCommon, xT, dw);
        if (!geo.ok || geo.stride != 1 || geo.dil != 1) return NOT_SUPPORT;
        const int C = xT->batch(), N = xT->channel(), H = xT->height(), W = xT->width();
        const int OC = dyT->length(0), OH = dyT->length(2), OW = dyT->length(3);
        const int KH = dw->height(), KW = dw->width();
        if (dyT->length(1) != N || OH != geo.kh || OW != geo.kw || dw->batch() != C || dw->channel() != OC) return NOT_SUPPORT;
        if (formatOf(dyT) != MNN_DATA_FORMAT_NCHW) return NOT_SUPPORT;
        mDesc = niti_conv_desc{N, C, H, W, OC, KH, KW, 1, geo.pad, 1};
        mOutCount = (size_t)OC * KH * KW * C;

        const int N4 = UP_DIV(N, 4) * 4;
        mX = (int8_t*)mScratch.get((size_t)N4 * H * W * C);
        mDy = (int8_t*)mScratch.get((size_t)N * OH * OW * OC);
        mAcc = (int32_t*)mScratch.get(mOutCount * 4);
        mDw = (int8_t*)mScratch.get(mOutCount);
        mMax = (uint32_t*)mScratch.get(256);
        if (mScratch.failed()) return OUT_OF_MEMORY;

        Gather gx; gx.src = dev(xT); gx.dst = mX;             // -> x [n][h*w][c]
        if (formatOf(xT) == MNN_DATA_FORMAT_NC4HW4) {         // buffer [N/4][C][HW][4], lane = n % 4
            gx.set(0, N4 / 4, (int64_t)C * H * W * 4); gx.set(1, 4, 1); gx.set(2, H * W, 4); gx.set(3, C, (int64_t)H * W * 4);
            gx.splitHi = 0; gx.splitLo = 1; gx.splitLimit = N;
        } else if (formatOf(xT) == MNN_DATA_FORMAT_NCHW) {
            gx.set(0, N, (int64_t)H * W); gx.set(1, H * W, 1); gx.set(2, C, (int64_t)N * H * W);
        } else {
            return NOT_SUPPORT;
        }
        mPre.emplace_back(gx);
        Gather gd; gd.src = dev(dyT); gd.dst = mDy;           // [OC][N][OH*OW] -> dy [n][oh*ow][oc]
        gd.set(0, N, (int64_t)OH * OW); gd.set(1, OH * OW, 1); gd.set(2, OC, (int64_t)N * OH * OW);
        mPre.emplace_back(gd);

        Gather go; go.src = mDw; go.dst = dev(dw);            // [oc][kh*kw][c] -> logical [C, OC, KH, KW]
        const int KK = KH * KW;
        if (formatOf(dw) == MNN_DATA_FORMAT_NC4HW4) {         // buffer [OC/4][C][KK][4]
            go.set(0, UP_DIV(OC, 4), (int64_t)4 * KK * C); go.set(1, C, 1); go.set(2, KK, C); go.set(3, 4, (int64_t)KK * C);
            go.splitHi = 0; go.splitLo = 3; go.splitLimit = OC;
        } else if (formatOf(dw) == MNN_DATA_FORMAT_NCHW) {
            go.set(0, C, 1); go.set(1, OC, (int64_t)KK * C); go.set(2, KK, C);
        } else {
            return NOT_SUPPORT;
        }
        mPost.emplace_back(go);

        mSlices = 1;
        if (mSplit) {
            // slice count of NITI_GradientSplitConv_Int8.cpp:209-222 (weights of one slice must fit 8192/(OH*OW))
            int align32 = OC % 32;
            if (align32 != 0) align32 = (OC / 32 + 1) * 32;
            const int high = 8192 / OW / OH;
            int realBatch = OC;
            while (align32 > high && realBatch > 1) { align32 /= 2; realBatch /= 2; }
            mSlices = realBatch > 0 ? OC / realBatch : 1;
            if (mSlices < 1 || OC % mSlices != 0 || mSlices > 64) return NOT_SUPPORT;
        }
        mScratch.releaseAll();
        return NO_ERROR;
    }

    virtual ErrorCode onExecute(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto& a = bn()->api(); auto ctx = bn()->ctx(); auto st = bn()->stream();
        auto code = runAll(mPre, "operand gather");
        if (code != NO_ERROR) return code;
        int rc = a.zero_u32(ctx, mMax, 64, st);
        if (rc == NITI_OK) rc = a.conv_dw_acc(ctx, &mDesc, mX, mDy, mAcc, mSplit ? nullptr : mMax, st);
        if (rc != NITI_OK) return check(rc, "niti_conv_dw_acc");
        if (!mSplit) {
            rc = a.rescale_grad(ctx, mAcc, mOutCount, mMax, 2, mDw, nullptr, st);          // NITI_GradientConv_Int8.cpp:274-296
        } else {
            const size_t per = mOutCount / mSlices;                                          // an OC slice is contiguous in [oc][kh][kw][c]
            for (int s = 0; s < mSlices && rc == NITI_OK; ++s) rc = a.absmax(ctx, mAcc + s * per, per, mMax + s, st);
            if (rc == NITI_OK) rc = a.rescale_grad_split(ctx, mAcc, per, mSlices, mMax, 3, mDw, st);
        }
        if (rc != NITI_OK) return check(rc, "niti_rescale_grad");
        return runAll(mPost, "result gather");
    }

private:
    const Convolution2DCommon* mCommon;
    bool mSplit;
    int mSlices = 1;
    niti_conv_desc mDesc;
    size_t mOutCount = 0;
    int8_t *mX = nullptr, *mDy = nullptr, *mDw = nullptr;
    int32_t* mAcc = nullptr;
    uint32_t* mMax = nullptr;
};

// ------------------------------------------------------------------------------------------------
// OpType_NITI_MatMul_Int8: {A [e, l], B [h, l]} -> {C [e, h]},  C = A * B^T, PSTO(acc, b - 3)
// (NITI_Matmul_Int8.cpp:140-231; emitted by source/geometry/NITI_GeometryConv2DBackPropFilter_Int8.cpp:78-126).
// The reference reads A and writes C as packed [channel/4][batch][1][4] blocks; B is row-major.
class B200NitiMatMul : public NitiExecution {
public:
    explicit B200NitiMatMul(Backend* bn) : NitiExecution(bn) {}

    virtual ErrorCode onResize(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto A = inputs[0], B = inputs[1], C = outputs[0];
        mPre.clear(); mPost.clear(); mScratch.clear();
        if (!isInt8(A) || !isInt8(B)) return NOT_SUPPORT;
        mE = A->length(0); mL = A->length(1); mH = C->length(1);
        if (B->length(1) != mL || B->length(0) != mH || C->length(0) != mE) return NOT_SUPPORT;
        if (mL % 4 != 0 || mH % 4 != 0) return NOT_SUPPORT;                                 // quirk Q6
        mA = (int8_t*)mScratch.get((size_t)mE * mL);
        mAcc = (int32_t*)mScratch.get((size_t)mE * mH * 4);
        mC = (int8_t*)mScratch.get((size_t)mE * mH);
        mMax = (uint32_t*)mScratch.get(256);
        if (mScratch.failed()) return OUT_OF_MEMORY;
        Gather ga; ga.src = dev(A); ga.dst = mA;              // [l/4][e][4] -> [e][l]
        ga.set(0, mE, 4); ga.set(1, mL / 4, (int64_t)mE * 4); ga.set(2, 4, 1);
        mPre.emplace_back(ga);
        Gather gc; gc.src = mC; gc.dst = dev(C);              // [e][h] -> [h/4][e][4]
        gc.set(0, mH / 4, 4); gc.set(1, mE, mH); gc.set(2, 4, 1);
        mPost.emplace_back(gc);
        mScratch.releaseAll();
        return NO_ERROR;
    }

    virtual ErrorCode onExecute(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto& a = bn()->api(); auto ctx = bn()->ctx(); auto st = bn()->stream();
        auto code = runAll(mPre, "operand gather");
        if (code != NO_ERROR) return code;
        int rc = a.zero_u32(ctx, mMax, 1, st);
        if (rc == NITI_OK) rc = a.matmul_nt_acc(ctx, mE, mH, mL, mA, dev(inputs[1]), mAcc, mMax, st);
        if (rc == NITI_OK) rc = a.rescale_grad(ctx, mAcc, (size_t)mE * mH, mMax, 3, mC, nullptr, st);
        if (rc != NITI_OK) return check(rc, "niti_matmul_nt");
        return runAll(mPost, "result gather");
    }

private:
    int mE = 0, mL = 0, mH = 0;
    int8_t *mA = nullptr, *mC = nullptr;
    int32_t* mAcc = nullptr;
    uint32_t* mMax = nullptr;
};

// ------------------------------------------------------------------------------------------------
// DSP-named op types.  The reference builds a hexagon_nn graph per op and ships it to the cDSP
// over FastRPC; operands are NHWC and weights [KH,KW,C,OC] (tools/train/source/nn/NN.cpp:1155-1157), the int8 values are
// re-biased to uint8 around the supernode and its int32 accumulators are requantised with round-to-nearest
// (niti_op_supernode_new.c:4614-4634).  Here the same contractions run on tcgen05 and niti_rescale_dsp applies that rule.
//   OpType_NITI_DSP_CONV_Int8    NITI_DSPConv_Int8::onExecute    source/backend/cpu/NITI_DSPConv_Int8.cpp:160-455
//   OpType_NITI_DSP_DECONV_Int8  NITI_DSPDeConv_Int8::onExecute  NITI_DSPDeConv_Int8.cpp:120-402 (a forward convolution of
//       dY with the rotated, transposed weights the grad rule hands it, grad/NITI_DSPConv_Int8_Grad.cpp:36-124)
//   {x [N,H,W,C], w [KH,KW,C,OC], int8 exp, int8 exp} -> {y [N,OH,OW,OC], exp = e0 + e1 + shift}
class B200NitiDspConv : public NitiExecution {
public:
    B200NitiDspConv(Backend* bn, const Convolution2DCommon* common) : NitiExecution(bn), mCommon(common) {}

    virtual ErrorCode onResize(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto x = inputs[0], w = inputs[1], y = outputs[0];
        mPre.clear(); mPost.clear(); mScratch.clear();
        if (!isInt8(x) || !isInt8(w) || formatOf(x) != MNN_DATA_FORMAT_NHWC || w->dimensions() != 4) return NOT_SUPPORT;
        auto geo = readGeometry(mCommon, x, y);
        if (!geo.ok) return NOT_SUPPORT;
        const int n = x->batch(), c = x->channel(), h = x->height(), wd = x->width();
        const int oc = y->channel(), oh = y->height(), ow = y->width();
        if (w->length(0) != geo.kh || w->length(1) != geo.kw || w->length(2) != c || w->length(3) != oc) return NOT_SUPPORT;
        mDesc = niti_conv_desc{n, c, h, wd, oc, geo.kh, geo.kw, geo.stride, geo.pad, geo.dil};
        mOutCount = (size_t)n * oh * ow * oc;
        const int kk = geo.kh * geo.kw;
        mW = (int8_t*)mScratch.get((size_t)oc * kk * c);
        mAcc = (int32_t*)mScratch.get(mOutCount * 4);
        mMax = (uint32_t*)mScratch.get(256);
        if (mScratch.failed()) return OUT_OF_MEMORY;
        Gather gw; gw.src = dev(w); gw.dst = mW;              // [taps][c][oc] -> [oc][taps][c]
        gw.set(0, oc, 1); gw.set(1, kk, (int64_t)c * oc); gw.set(2, c, oc);
        mPre.emplace_back(gw);
        mScratch.releaseAll();
        return NO_ERROR;
    }

    virtual ErrorCode onExecute(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto& a = bn()->api(); auto ctx = bn()->ctx(); auto st = bn()->stream();
        auto code = runAll(mPre, "weight gather");
        if (code != NO_ERROR) return code;
        int rc = a.zero_u32(ctx, mMax, 1, st);
        if (rc == NITI_OK) rc = a.conv_fwd_acc(ctx, &mDesc, dev(inputs[0]), mW, mAcc, mMax, st);   // NHWC in place: no copy
        if (rc != NITI_OK) return check(rc, "niti_conv_fwd_acc");
        const int8_t* e0 = inputs.size() > 2 ? dev(inputs[2]) : nullptr;
        const int8_t* e1 = inputs.size() > 3 ? dev(inputs[3]) : nullptr;
        int8_t* eo = outputs.size() > 1 ? dev(outputs[1]) : nullptr;
        rc = a.rescale_dsp(ctx, mAcc, mOutCount, mMax, 1, dev(outputs[0]), e0, e1, eo, st);
        return rc == NITI_OK ? NO_ERROR : check(rc, "niti_rescale_dsp");
    }

private:
    const Convolution2DCommon* mCommon;
    niti_conv_desc mDesc;
    size_t mOutCount = 0;
    int8_t* mW = nullptr;
    int32_t* mAcc = nullptr;
    uint32_t* mMax = nullptr;
};

// OpType_NITI_DSP_TRANSPOSEGRADIENT_CONV_Int8 / _PARALLEL_GRADIENTCONV / _GRADIENT_SPLITBatchCONV / _GRADIENTCONV:
//   {xT [C,H,W,N] (x transposed {3,1,2,0}), dY [N,OH,OW,OC] (zero-inserted for stride-2 layers), exp, exp}
//   -> {dW laid out [KH,KW,C,OC], exp}     (grad/NITI_DSPConv_Int8_Grad.cpp:126-222)
// The reference convolves xT with dY transposed to [OH,OW,N,OC] as the filter and transposes the result back
// (NITI_DSPTransposeGradientConv_Int8.cpp:262-270,324-330); that is the weight gradient niti_conv_dw_acc computes from
// the natural operands.  The int8 result is divided by 32 (TRANSPOSEGRADIENT, :430-434) or 16 (the other variants,
// NITI_DSPParallelGradientConv_Int8.cpp:432, NITI_DSPGradientSplitBatchConv_Int8.cpp:583, NITI_DSPGradientConv_Int8.cpp:430).
// The batch / output-channel splitting of the PARALLEL and SPLITBatch variants exists to fit the cDSP's VTCM and to
// co-run with the CPU; one B200 launch covers the whole contraction, so they share this execution.
class B200NitiDspGradientConv : public NitiExecution {
public:
    B200NitiDspGradientConv(Backend* bn, const Convolution2DCommon* common, int postDiv, bool transposedX)
        : NitiExecution(bn), mCommon(common), mPostDiv(postDiv), mTransposedX(transposedX) {}

    virtual ErrorCode onResize(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto xT = inputs[0], dy = inputs[1], dw = outputs[0];
        mPre.clear(); mPost.clear(); mScratch.clear();
        if (!isInt8(xT) || !isInt8(dy) || xT->dimensions() != 4 || dy->dimensions() != 4) { MNN_ERROR("[niti_b200] DSP gradient conv: unsupported operands (check %d)\n", 1); return NOT_SUPPORT; }
        if (mCommon->strideX() != 1 || mCommon->strideY() != 1 || mCommon->dilateX() != 1 || mCommon->dilateY() != 1) { MNN_ERROR("[niti_b200] DSP gradient conv: unsupported operands (check %d)\n", 2); return NOT_SUPPORT; }
        if (mCommon->padX() != mCommon->padY() || mCommon->padX() < 0 || mCommon->padMode() == PadMode_SAME) { MNN_ERROR("[niti_b200] DSP gradient conv: unsupported operands (check %d)\n", 3); return NOT_SUPPORT; }
        // TRANSPOSEGRADIENT / SPLITBatch take x transposed {3,1,2,0}; PARALLEL / DSP_GRADIENTCONV take it as it is — the
        // reference's shape rules say which: output dim 2 = input dim 0 resp. input dim 3
        // (source/shape/ShapeNITI_Conv_Int8.cpp:220-224 and :312-315, registered at :527-531)
        const int C = xT->length(mTransposedX ? 0 : 3), H = xT->length(1), W = xT->length(2), N = xT->length(mTransposedX ? 3 : 0);
        const int OH = dy->length(1), OW = dy->length(2), OC = dy->length(3), pad = mCommon->padX();
        const int KH = H + 2 * pad - OH + 1, KW = W + 2 * pad - OW + 1;
        if (dy->length(0) != N || KH < 1 || KW < 1) { MNN_ERROR("[niti_b200] DSP gradient conv: unsupported operands (check %d)\n", 4); return NOT_SUPPORT; }
        mCount = (size_t)OC * KH * KW * C;
        if ((size_t)dw->elementSize() != mCount) { MNN_ERROR("[niti_b200] DSP gradient conv: unsupported operands (check %d)\n", 5); return NOT_SUPPORT; }
        mDesc = niti_conv_desc{N, C, H, W, OC, KH, KW, 1, pad, 1};
        mX = (int8_t*)mScratch.get((size_t)N * H * W * C);
        mAcc = (int32_t*)mScratch.get(mCount * 4);
        mDw = (int8_t*)mScratch.get(mCount);
        mMax = (uint32_t*)mScratch.get(256);
        if (mScratch.failed()) return OUT_OF_MEMORY;
        if (mTransposedX) {
            Gather gx; gx.src = dev(xT); gx.dst = mX;         // [C][H*W][N] -> [N][H*W][C]
            gx.set(0, N, 1); gx.set(1, H * W, N); gx.set(2, C, (int64_t)H * W * N);
            mPre.emplace_back(gx);
        }
        Gather go; go.src = mDw; go.dst = dev(dw);            // [OC][taps][C] -> [taps][C][OC]
        go.set(0, KH * KW, C); go.set(1, C, 1); go.set(2, OC, (int64_t)KH * KW * C);
        mPost.emplace_back(go);
        mScratch.releaseAll();
        return NO_ERROR;
    }

    virtual ErrorCode onExecute(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs) override {
        auto& a = bn()->api(); auto ctx = bn()->ctx(); auto st = bn()->stream();
        auto code = runAll(mPre, "operand gather");
        if (code != NO_ERROR) return code;
        int rc = a.zero_u32(ctx, mMax, 1, st);
        if (rc == NITI_OK) rc = a.conv_dw_acc(ctx, &mDesc, mTransposedX ? mX : dev(inputs[0]), dev(inputs[1]), mAcc, mMax, st);
        if (rc != NITI_OK) return check(rc, "niti_conv_dw_acc");
        const int8_t* e0 = inputs.size() > 2 ? dev(inputs[2]) : nullptr;
        const int8_t* e1 = inputs.size() > 3 ? dev(inputs[3]) : nullptr;
        int8_t* eo = outputs.size() > 1 ? dev(outputs[1]) : nullptr;
        rc = a.rescale_dsp(ctx, mAcc, mCount, mMax, mPostDiv, mDw, e0, e1, eo, st);
        if (rc != NITI_OK) return check(rc, "niti_rescale_dsp");
        return runAll(mPost, "result gather");
    }

private:
    const Convolution2DCommon* mCommon;
    int mPostDiv;
    bool mTransposedX;
    niti_conv_desc mDesc;
    size_t mCount = 0;
    int8_t *mX = nullptr, *mDw = nullptr;
    int32_t* mAcc = nullptr;
    uint32_t* mMax = nullptr;
};

// ------------------------------------------------------------------------------------------------
class ConvCreator : public B200Backend::Creator {
public:
    virtual Execution* onCreate(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs,
                                const MNN::Op* op, Backend* backend) const override {
        auto conv = op->main_as_NITI_CONV_Int8();
        if (nullptr == conv || nullptr == conv->common()) return nullptr;
        switch (op->type()) {
            case OpType_NITI_CONV_Int8:
                if (inputs.size() < 4 || outputs.size() < 2) return nullptr;
                return new B200NitiConv(backend, conv->common(), false);
            case OpType_NITI_DeCONV_Int8:
                if (inputs.size() < 2) return nullptr;
                return new B200NitiConv(backend, conv->common(), true);
            case OpType_NITI_GradientCONV_Int8:
                if (inputs.size() < 2) return nullptr;
                return new B200NitiGradientConv(backend, conv->common(), false);
            case OpType_NITI_GradientSplitCONV_Int8:
                if (inputs.size() < 2) return nullptr;
                return new B200NitiGradientConv(backend, conv->common(), true);
            case OpType_NITI_DSP_CONV_Int8:
            case OpType_NITI_DSP_DECONV_Int8:
                if (inputs.size() < 2) return nullptr;
                return new B200NitiDspConv(backend, conv->common());
            case OpType_NITI_DSP_TRANSPOSEGRADIENT_CONV_Int8:
                if (inputs.size() < 2) return nullptr;
                return new B200NitiDspGradientConv(backend, conv->common(), 32, true);
            case OpType_NITI_DSP_GRADIENT_SPLITBatchCONV_Int8:
                if (inputs.size() < 2) return nullptr;
                return new B200NitiDspGradientConv(backend, conv->common(), 16, true);
            case OpType_NITI_DSP_GRADIENTCONV_Int8:
            case OpType_NITI_DSP_PARALLEL_GRADIENTCONV_Int8:
            // OpType_NITI_DSP_MATMUL_GRADIENT_Int8 — the op BASELINE.json's north star names.  Operand contract read off the
            // execution and its size computer: shape rule NITI_DSPGradientCONVSizeComputer, the SAME one as DSP_GRADIENTCONV
            // (source/shape/ShapeNITI_Conv_Int8.cpp:146-236, registered :528); inputs[0] = x NHWC [N,H,W,C], inputs[1] = dY
            // NHWC [N,OH,OW,OC] (NITI_DSPMatmulGradientConv_Int8.cpp:110-126), kernelX/Y = OW/OH, zero-padded by padX (:153);
            // the hexagon graph is pad -> im2col -> transpose {3,1,2,0} -> [C*KH*KW, N*OH*OW] x dY as [N*OH*OW, OC]
            // (QuantizedMatMul_8x8to32) -> requantise -> transpose to [KH,KW,C,OC], then the int8 result is divided by 16
            // (:548-550).  That is the dW contraction with the DSP rounding and post-division 16.
            case OpType_NITI_DSP_MATMUL_GRADIENT_Int8:
            // OpType_NITI_DSP_MATMUL_Int8 — same size computer (ShapeNITI_Conv_Int8.cpp:532), same operands, same /16
            // (NITI_DSPMatmul_Int8.cpp:297-302).  The reference execution is unfinished: its graph is the bare
            // QuantizedMatMul_8x8to32 of scratch tensors matmulA [C*KH*KW, OH*OW*N] x matmulB [OH*OW*N, OC] (:72-87,193-262)
            // that are never filled from the inputs, and outputs[0] is never written by the graph.  What IS defined — the
            // shape rule, the matmul dimensions (exactly the im2col form of the dW contraction) and the /16 — is what
            // MATMUL_GRADIENT computes, so both op types share this execution.
            case OpType_NITI_DSP_MATMUL_Int8:
                if (inputs.size() < 2) return nullptr;
                return new B200NitiDspGradientConv(backend, conv->common(), 16, false);
            default:
                return nullptr;
        }
    }
};

class MatMulCreator : public B200Backend::Creator {
public:
    virtual Execution* onCreate(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs,
                                const MNN::Op* op, Backend* backend) const override {
        if (inputs.size() < 2) return nullptr;
        return new B200NitiMatMul(backend);
    }
};

typedef ConvCreator ConvFwdCreator;
typedef ConvCreator ConvDxCreator;
typedef ConvCreator ConvDwCreator;
typedef ConvCreator ConvDwSplitCreator;
REGISTER_B200_OP_CREATOR(ConvFwdCreator, OpType_NITI_CONV_Int8);
REGISTER_B200_OP_CREATOR(ConvDxCreator, OpType_NITI_DeCONV_Int8);
REGISTER_B200_OP_CREATOR(ConvDwCreator, OpType_NITI_GradientCONV_Int8);
REGISTER_B200_OP_CREATOR(ConvDwSplitCreator, OpType_NITI_GradientSplitCONV_Int8);
REGISTER_B200_OP_CREATOR(MatMulCreator, OpType_NITI_MatMul_Int8);
typedef ConvCreator DspConvCreator;
typedef ConvCreator DspDeconvCreator;
typedef ConvCreator DspDwCreator;
typedef ConvCreator DspDwTransposeCreator;
typedef ConvCreator DspDwParallelCreator;
typedef ConvCreator DspDwSplitBatchCreator;
REGISTER_B200_OP_CREATOR(DspConvCreator, OpType_NITI_DSP_CONV_Int8);
REGISTER_B200_OP_CREATOR(DspDeconvCreator, OpType_NITI_DSP_DECONV_Int8);
REGISTER_B200_OP_CREATOR(DspDwCreator, OpType_NITI_DSP_GRADIENTCONV_Int8);
REGISTER_B200_OP_CREATOR(DspDwTransposeCreator, OpType_NITI_DSP_TRANSPOSEGRADIENT_CONV_Int8);
REGISTER_B200_OP_CREATOR(DspDwParallelCreator, OpType_NITI_DSP_PARALLEL_GRADIENTCONV_Int8);
REGISTER_B200_OP_CREATOR(DspDwSplitBatchCreator, OpType_NITI_DSP_GRADIENT_SPLITBatchCONV_Int8);
typedef ConvCreator DspMatmulGradientCreator;
typedef ConvCreator DspMatmulCreator;
REGISTER_B200_OP_CREATOR(DspMatmulGradientCreator, OpType_NITI_DSP_MATMUL_GRADIENT_Int8);
REGISTER_B200_OP_CREATOR(DspMatmulCreator, OpType_NITI_DSP_MATMUL_Int8);

} // namespace
} // namespace B200
} // namespace MNN
