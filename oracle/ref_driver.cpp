// ref_driver.cpp — TEST INFRASTRUCTURE ONLY.
//
// Drives the UNMODIFIED reference (oracle/_ref/libMNN_ref.so + libMNNTrain_ref.so, built from
// /root/reference by oracle/Makefile.ref) op by op through its own public express API and its own
// autodiff rule, with injected int8 tensors, so the C restatement in niti_oracle.c can be pinned
// against the real thing and golden vectors can be generated (tests/golden/make_golden.py).
// Exposes a tiny C ABI for ctypes.  All tensors cross this boundary as NCHW / OIHW int8.
//
// Reference entry points exercised:
//   _NITI_Conv_Int8            express/NeuralNetWorkOp.cpp:1839-1888   -> NITI_Conv_Int8::onExecute
//   NITI_Conv_Int8_Grad::onGrad tools/train/source/grad/NITI_Conv_Int8_Grad.cpp:16-197
//                                                 -> NITI_DeConv_Int8 / NITI_GradientConv_Int8
//   int8 Binary SUB (NITI_SGD.hpp:49-52)          -> NITI_CPUBinary_Int8
//   _NITI_Relu_Int8 / _NITI_Maxpool_Int8 / OpGrad of those / NITI_LOSS_Grad
//   whole training steps: NN::NITI_Conv_Int8 modules + _NITI_LOSS_SUM + NITI_SGD::step, the op sequence of
//   demo/mnistTrain.cpp:131-182 and demo/MnistUtils.cpp:79-147 with injected weights, data and labels
//   (ref_net_train); ref_use_backend switches the executor to another registered forward type, which is
//   how tests/test_gpu_mnn_backend.py runs the SAME reference graph on the B200 plug-in (mnn_backend/).
// The reference reads DSP.txt / parallel.txt from the cwd with unchecked fopen
// (hexagon_implementation.cpp:33-39, CPUBackend.cpp:155-157): ref_init() writes them (2 and 0).
#include <MNN/expr/Expr.hpp>
#include <MNN/expr/ExprCreator.hpp>
#include <MNN/expr/Executor.hpp>
#include "OpGrad.hpp"
#include "NN.hpp"
#include "core/Backend.hpp"
#include "NITI_SGD.hpp"
#include "MNN_generated.h"
#include <MNN/expr/Module.hpp>
#include <dlfcn.h>
#include <chrono>
#include <memory>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <unistd.h>

using namespace MNN;
using namespace MNN::Express;

namespace {
VARP constI8(const int8_t* p, std::vector<int> dims, Dimensionformat fmt = NCHW) {
    return _Const(p, dims, fmt, halide_type_of<int8_t>());
}
void copyOut(VARP v, int8_t* dst) {
    auto info = v->getInfo();
    auto ptr  = v->readMap<int8_t>();
    ::memcpy(dst, ptr, info->size);
}
struct Shape { int n, c, h, w, oc, kh, kw, stride, pad, dil; };
std::vector<VARP> fwd(const Shape& s, VARP xc, VARP w, int ex, int ew) {
    auto ascale = _Scalar<int8_t>((int8_t)ex);
    auto wscale = _Scalar<int8_t>((int8_t)ew);
    return _NITI_Conv_Int8(w, ascale, wscale, xc, CAFFE, {s.stride, s.stride}, {s.dil, s.dil}, 1,
                           {s.pad, s.pad}, 8);
}
} // namespace

extern "C" {

int ref_init(const char* workdir, int threads) {
    if (workdir && ::chdir(workdir) != 0) return -1;
    FILE* f = ::fopen("DSP.txt", "w");      if (!f) return -2; ::fputs("2", f); ::fclose(f);
    f = ::fopen("parallel.txt", "w");       if (!f) return -3; ::fputs("0", f); ::fclose(f);
    BackendConfig config;
    Executor::getGlobalExecutor()->setGlobalExecutorConfig(MNN_FORWARD_CPU, config, threads);
    return 0;
}

// y [n,oc,oh,ow] NCHW; returns exp_out
int ref_conv_fwd(const int* shp, const int8_t* x, const int8_t* w, int ex, int ew, int8_t* y) {
    Shape s; ::memcpy(&s, shp, sizeof(s));
    auto xv = _Convert(constI8(x, {s.n, s.c, s.h, s.w}), NC4HW4);
    auto wv = constI8(w, {s.oc, s.c, s.kh, s.kw});
    auto res = fwd(s, xv, wv, ex, ew);
    auto yv  = _Convert(res[0], NCHW);
    copyOut(yv, y);
    return (int)res[1]->readMap<int8_t>()[0];
}

// dx [n,c,h,w] NCHW, dw [oc,c,kh,kw]; both through the reference's own grad rule.
int ref_conv_bwd(const int* shp, const int8_t* x, const int8_t* w, const int8_t* dy, int ex, int ew,
                 int8_t* dx, int8_t* dw) {
    Shape s; ::memcpy(&s, shp, sizeof(s));
    auto xv = _Convert(constI8(x, {s.n, s.c, s.h, s.w}), NC4HW4);
    auto wv = constI8(w, {s.oc, s.c, s.kh, s.kw});
    auto res = fwd(s, xv, wv, ex, ew);
    auto info = res[0]->getInfo();
    if (!info) return -1;
    int oh = info->dim[2], ow = info->dim[3];
    auto dyv = _Convert(constI8(dy, {s.n, s.oc, oh, ow}), NC4HW4);
    auto rule = OpGrad::get(OpType_NITI_CONV_Int8);
    if (!rule) return -2;
    auto grads = rule->onGrad(res[0]->expr().first, {dyv, nullptr});
    if (grads.size() < 2 || grads[0].get() == nullptr || grads[1].get() == nullptr) return -3;
    if (dx) copyOut(_Convert(grads[0], NCHW), dx);
    if (dw) copyOut(_Convert(grads[1], NCHW), dw);
    return 0;
}

// w <- w - dw through the reference's int8 binary op
int ref_sgd(const int8_t* w, const int8_t* dw, int n, int8_t* out) {
    auto a = constI8(w, {n}), b = constI8(dw, {n});
    auto c = a - b;
    copyOut(c, out);
    return 0;
}

int ref_relu(const int8_t* x, int n, int c, int h, int w, int8_t* y) {
    auto xv = _Convert(constI8(x, {n, c, h, w}), NC4HW4);
    copyOut(_Convert(_NITI_Relu_Int8(xv), NCHW), y);
    return 0;
}
int ref_relu_grad(const int8_t* x, const int8_t* dy, int n, int c, int h, int w, int8_t* dx) {
    auto xv  = _Convert(constI8(x, {n, c, h, w}), NC4HW4);
    auto dyv = _Convert(constI8(dy, {n, c, h, w}), NC4HW4);
    copyOut(_Convert(_NITI_ReluGrad_Int8(xv, dyv), NCHW), dx);
    return 0;
}
int ref_maxpool(const int8_t* x, int n, int c, int h, int w, int k, int st, int8_t* y) {
    auto xv = _Convert(constI8(x, {n, c, h, w}), NC4HW4);
    auto res = _NITI_Maxpool_Int8(xv, _Scalar<int8_t>(0), {k, k}, {st, st});
    copyOut(_Convert(res[0], NCHW), y);
    return 0;
}
int ref_maxpool_grad(const int8_t* x, const int8_t* dy, int n, int c, int h, int w, int k, int st, int8_t* dx) {
    auto xv = _Convert(constI8(x, {n, c, h, w}), NC4HW4);
    auto res = _NITI_Maxpool_Int8(xv, _Scalar<int8_t>(0), {k, k}, {st, st});
    auto info = res[0]->getInfo();
    // the pool's output (and so its gradient) is NHWC in the reference's graphs (express/NeuralNetWorkOp.cpp:1911)
    auto dyv = _Convert(constI8(dy, {n, c, info->dim[1], info->dim[2]}), NHWC);
    auto rule = OpGrad::get(OpType_NITI_Maxpool_Int8);
    if (!rule) return -2;
    auto grads = rule->onGrad(res[0]->expr().first, {dyv, nullptr});
    if (grads.empty() || grads[0].get() == nullptr) return -3;
    copyOut(_Convert(grads[0], NCHW), dx);
    return 0;
}
// integer softmax-CE gradient: logits [b,classes,1,1], onehot int32 [b,classes]
int ref_loss_grad(const int8_t* logits, int ascale, const int* onehot, int b, int classes, int8_t* grad) {
    auto lv = constI8(logits, {b, classes, 1, 1});
    auto tv = _Const(onehot, {b, classes}, NCHW, halide_type_of<int32_t>());
    auto loss = _NITI_LOSS_SUM(lv, _Scalar<int8_t>((int8_t)ascale), tv, b, classes, 1, 1);
    auto rule = OpGrad::get(loss->expr().first->get()->type());
    if (!rule) return -2;
    auto one = _Const(1.0f);
    auto grads = rule->onGrad(loss->expr().first, {one});
    if (grads.empty() || grads[0].get() == nullptr) return -3;
    copyOut(grads[0], grad);
    return 0;
}
// the loss scalar itself: OpType_NITI_LOSS_Int8 -> NITI_CPULoss_Int8::onExecute (source/backend/cpu/NITI_CPULoss_Int8.cpp:68-131)
int ref_loss_value(const int8_t* logits, int ascale, const int* onehot, int b, int classes, float* out) {
    auto lv = constI8(logits, {b, classes, 1, 1});
    auto tv = _Const(onehot, {b, classes}, NCHW, halide_type_of<int32_t>());
    auto loss = _NITI_LOSS_SUM(lv, _Scalar<int8_t>((int8_t)ascale), tv, b, classes, 1, 1);
    auto lp = loss->readMap<float>();
    if (lp == nullptr) return -2;
    *out = lp[0];
    return 0;
}


// ---- backend selection ----------------------------------------------------------------------------
// dlopen an MNN backend plug-in (its static registrar calls MNNInsertExtraRuntimeCreator) and make
// `forwardType` the executor's backend.  Returns 0 when the executor really switched (the reference's
// Executor::setGlobalExecutorConfig silently keeps the old runtime when no creator answers,
// express/Executor.cpp:81-86, so the switch is verified through the runtime's compiler/forward type).
// One DSP-named op through the reference's express API: Expr::create with the op type, a
// Convolution2DCommon and NHWC operands, exactly the form grad/NITI_DSPConv_Int8_Grad.cpp:36-222 and
// express/NeuralNetWorkOp.cpp:2031-2100 emit.  On the reference CPU backend these executions need the Hexagon FastRPC
// library (not runnable here); tests/test_gpu_mnn_backend.py runs them on the B200 plug-in only.
// kind: 0 DSP_CONV, 1 DSP_DECONV, 2 TRANSPOSEGRADIENT_CONV, 3 PARALLEL_GRADIENTCONV, 4 GRADIENT_SPLITBatchCONV, 5 DSP_GRADIENTCONV,
//       6 DSP_MATMUL_GRADIENT (NITI_DSPMatmulGradientConv_Int8.cpp:105-553), 7 DSP_MATMUL (NITI_DSPMatmul_Int8.cpp:142-305)
// geo = {kernelY, kernelX, stride, pad, inputCount, outputCount}; returns the exponent output, odims = output dims.
int ref_dsp_op(int kind, const int* geo, const int8_t* a, const int* adims, const int8_t* b, const int* bdims, int e0, int e1,
               int8_t* out, int* odims) {
    static const OpType types[] = {OpType_NITI_DSP_CONV_Int8, OpType_NITI_DSP_DECONV_Int8,
                                   OpType_NITI_DSP_TRANSPOSEGRADIENT_CONV_Int8, OpType_NITI_DSP_PARALLEL_GRADIENTCONV_Int8,
                                   OpType_NITI_DSP_GRADIENT_SPLITBatchCONV_Int8, OpType_NITI_DSP_GRADIENTCONV_Int8,
                                   OpType_NITI_DSP_MATMUL_GRADIENT_Int8, OpType_NITI_DSP_MATMUL_Int8};
    if (kind < 0 || kind > 7) return -1000;
    std::unique_ptr<OpT> op(new OpT);
    op->type = types[kind];
    op->main.type = OpParameter_NITI_CONV_Int8;
    auto conv = new NITI_CONV_Int8T;
    conv->common.reset(new Convolution2DCommonT);
    conv->common->kernelY = geo[0]; conv->common->kernelX = geo[1];
    conv->common->strideX = conv->common->strideY = geo[2];
    conv->common->padX = conv->common->padY = geo[3];
    conv->common->padMode = PadMode_CAFFE;
    conv->common->dilateX = conv->common->dilateY = 1;
    conv->common->group = 1;
    conv->common->inputCount = geo[4]; conv->common->outputCount = geo[5];
    op->main.value = conv;
    auto av = constI8(a, {adims[0], adims[1], adims[2], adims[3]}, NHWC);
    auto bv = constI8(b, {bdims[0], bdims[1], bdims[2], bdims[3]}, NHWC);
    auto expr = Expr::create(std::move(op), {av, bv, _Scalar<int8_t>((int8_t)e0), _Scalar<int8_t>((int8_t)e1)}, 2);
    auto y = Variable::create(expr, 0), e = Variable::create(expr, 1);
    auto info = y->getInfo();
    if (!info || info->dim.size() != 4) return -1001;
    for (int i = 0; i < 4; ++i) odims[i] = info->dim[i];
    if (::getenv("REF_DSP_SHAPE_ONLY")) return 0;         // shape inference only (the CPU executions need the Hexagon library)
    auto ptr = y->readMap<int8_t>();
    if (!ptr) return -1002;
    ::memcpy(out, ptr, info->size);
    auto ep = e->readMap<int8_t>();
    return ep ? (int)ep[0] : -1003;
}

// OpType_NITI_MatMul_Int8 exactly as source/geometry/NITI_GeometryConv2DBackPropFilter_Int8.cpp:78-126 emits it:
// MatMul parameter with transposeB = true, inputs {A [e, l], B [h, l]} as 2-d int8 tensors, one output C [e, h].
// The buffers cross this boundary RAW (the execution reads A and writes C through its own packed [channel/4][batch][4]
// addressing, NITI_Matmul_Int8.cpp:186-208; tests restate that addressing).
int ref_matmul(int e, int l, int h, const int8_t* a, const int8_t* b, int8_t* c) {
    std::unique_ptr<OpT> op(new OpT);
    op->type = OpType_NITI_MatMul_Int8;
    op->main.type = OpParameter_MatMul;
    op->main.value = new MatMulT;
    op->main.AsMatMul()->transposeA = false;
    op->main.AsMatMul()->transposeB = true;
    auto av = constI8(a, {e, l});
    auto bv = constI8(b, {h, l});
    auto expr = Expr::create(std::move(op), {av, bv}, 1);
    auto y = Variable::create(expr, 0);
    auto info = y->getInfo();
    if (!info || info->dim.size() != 2 || info->dim[0] != e || info->dim[1] != h) return -1;
    auto ptr = y->readMap<int8_t>();
    if (!ptr) return -2;
    ::memcpy(c, ptr, (size_t)e * h);
    return 0;
}

// dW through OpType_NITI_GradientCONV_Int8 (split = 0) or OpType_NITI_GradientSplitCONV_Int8 (split = 1) with the operand
// forms of the reference's own grad rule, stride-1 branch (tools/train/source/grad/NITI_Conv_Int8_Grad.cpp:167-190):
// {xT [C,N,H,W] NC4HW4, dyT [OC,N,OH,OW] NCHW, ascale, wscale}, kernelX/Y = OW/OH, inputCount = N, outputCount = OC;
// result [C,OC,KH,KW] -> NCHW -> transpose {1,0,2,3} = dw [OC,C,KH,KW].  The split op type has no producer in the tree
// (SURVEY.md 8 a5); this is the only way to execute NITI_GradientSplitConv_Int8.cpp:195-386.
int ref_grad_conv(int split, const int* shp, const int8_t* x, const int8_t* dy, int8_t* dw) {
    Shape s; ::memcpy(&s, shp, sizeof(s));
    if (s.stride != 1 || s.dil != 1) return -1;
    const int oh = s.h + 2 * s.pad - s.kh + 1, ow = s.w + 2 * s.pad - s.kw + 1;
    std::unique_ptr<OpT> op(new OpT);
    op->type = split ? OpType_NITI_GradientSplitCONV_Int8 : OpType_NITI_GradientCONV_Int8;
    op->main.type = OpParameter_NITI_CONV_Int8;
    auto conv = new NITI_CONV_Int8T;
    conv->common.reset(new Convolution2DCommonT);
    conv->common->kernelX = ow; conv->common->kernelY = oh;
    conv->common->strideX = conv->common->strideY = 1;
    conv->common->dilateX = conv->common->dilateY = 1;
    conv->common->padX = conv->common->padY = s.pad;
    conv->common->padMode = PadMode_CAFFE;
    conv->common->group = 1;
    conv->common->inputCount = s.n; conv->common->outputCount = s.oc;
    op->main.value = conv;
    auto xv  = constI8(x, {s.n, s.c, s.h, s.w});
    auto dyv = constI8(dy, {s.n, s.oc, oh, ow});
    auto xT  = _Convert(_NITI_Transpose_INT8(xv, {1, 0, 2, 3}), NC4HW4);
    auto dyT = _NITI_Transpose_INT8(dyv, {1, 0, 2, 3});
    auto expr = Expr::create(std::move(op), {xT, dyT, _Scalar<int8_t>(-7), _Scalar<int8_t>(-7)}, 2);
    auto ret = Variable::create(expr, 0);
    auto out = _NITI_Transpose_INT8(_Convert(ret, NCHW), {1, 0, 2, 3});
    auto info = out->getInfo();
    if (!info || info->dim.size() != 4) return -2;
    if (info->dim[0] != s.oc || info->dim[1] != s.c || info->dim[2] != s.kh || info->dim[3] != s.kw) return -3;
    copyOut(out, dw);
    return 0;
}

int ref_use_backend(int forwardType, const char* plugin, int threads) {
    if (plugin && plugin[0]) {
        void* h = ::dlopen(plugin, RTLD_NOW | RTLD_GLOBAL);
        if (!h) { ::fprintf(stderr, "ref_use_backend: %s\n", ::dlerror()); return -1; }
    }
    if (forwardType != (int)MNN_FORWARD_CPU && MNNGetExtraRuntimeCreator((MNNForwardType)forwardType) == nullptr) return -2;
    BackendConfig config;
    Executor::getGlobalExecutor()->setGlobalExecutorConfig((MNNForwardType)forwardType, config, threads);
    return 0;
}

} // extern "C"

// ---- whole training steps through the reference's Module / OpGrad / NITI_SGD ------------------------
// layer list: 5 ints per layer {kind, a, b, c, d}: 0 conv(oc,k,stride,pad) 1 relu 2 maxpool(k,s) 3 flatten
namespace {
struct LayerSpec { int kind, a, b, c, d; };
class SpecNet : public Module {
public:
    SpecNet(const std::vector<LayerSpec>& layers, int inChannels, const int* chw, const int8_t* weights, const int8_t* wscales) : mLayers(layers) {
        int c = inChannels, h = chw[1], w = chw[2];
        size_t off = 0; int wi = 0;
        std::vector<std::shared_ptr<Module>> children;
        for (auto& L : mLayers) {
            if (L.kind == 0) {
                NN::NITI_Conv_Int8_Parameters p;
                p.option.kernelSize = {L.b, L.b}; p.option.channel = {c, L.a};
                p.option.stride = {L.c, L.c}; p.option.pads = {L.d, L.d}; p.option.padMode = Express::CAFFE;
                p.weight = _Const(weights + off, {L.a, c, L.b, L.b}, NCHW, halide_type_of<int8_t>());
                p.weight.fix(VARP::TRAINABLE);
                p.wscale = wscales[wi]; p.group = 1; p.name = "conv" + std::to_string(wi);
                off += (size_t)L.a * c * L.b * L.b; ++wi;
                std::shared_ptr<Module> m(NN::NITI_Conv_Int8(p, 8, false));
                mConvs.emplace_back(m); children.emplace_back(m);
                h = (h + 2 * L.d - L.b) / L.c + 1; w = (w + 2 * L.d - L.b) / L.c + 1; c = L.a;
            } else if (L.kind == 2) {
                h = (h - L.a) / L.b + 1; w = (w - L.a) / L.b + 1;
            } else if (L.kind == 3) {
                c = c * h * w; h = w = 1;
            }
        }
        registerModel(children);
    }
    virtual std::vector<VARP> onForward(const std::vector<VARP>& inputs) override {
        std::vector<VARP> x = inputs;
        int ci = 0;
        for (auto& L : mLayers) {
            switch (L.kind) {
                case 0: x = mConvs[ci++]->multioutput_forward({x[0], x[1]}); break;
                case 1: x[0] = _NITI_Relu_Int8(x[0]); break;
                case 2: x = _NITI_Maxpool_Int8(x[0], x[1], {L.a, L.a}, {L.b, L.b}); break;
                case 3: x[0] = _Reshape(_Convert(x[0], NCHW), {0, -1, 1, 1}); break;
            }
        }
        x[0] = _Reshape(_Convert(x[0], NCHW), {0, -1});
        return x;
    }
    std::vector<std::shared_ptr<Module>> mConvs;
private:
    std::vector<LayerSpec> mLayers;
};
} // namespace

extern "C" {

// Runs `steps` iterations of: forward, _NITI_LOSS_SUM, NITI_SGD::step (OpGrad over the whole graph, int8 SUB update).
//   x       steps x [n, c, h, w] int8 NCHW, ascale the input exponent; onehot steps x [n, tclasses] int32
//   weights concatenated OIHW int8 of the conv layers in order (in), wscales one per conv layer
//   out: logits steps x [n, classes] int8, logit exponents [steps], losses [steps] float, final weights (same packing),
//        ms_per_step [steps] wall time of each iteration
int ref_net_train(const int* spec, int nlayers, int n, const int* chw, int ascale, int steps, const int8_t* x,
                  const int* onehot, int tclasses, const int8_t* weights, const int8_t* wscales,
                  int8_t* logits_out, int8_t* elogits_out, float* loss_out, int8_t* weights_out, float* ms_out) {
    std::vector<LayerSpec> layers(nlayers);
    for (int i = 0; i < nlayers; ++i) layers[i] = LayerSpec{spec[5 * i], spec[5 * i + 1], spec[5 * i + 2], spec[5 * i + 3], spec[5 * i + 4]};
    std::shared_ptr<Module> model(new SpecNet(layers, chw[0], chw, weights, wscales));
    model->setIsTraining(true);
    std::shared_ptr<MNN::Train::SGD> sgd(new MNN::Train::NITI_SGD(model));
    const size_t xs = (size_t)n * chw[0] * chw[1] * chw[2];
    int classes = 0;
    for (int s = 0; s < steps; ++s) {
        auto t0 = std::chrono::steady_clock::now();
        auto xv = _Const(x + s * xs, {n, chw[0], chw[1], chw[2]}, NCHW, halide_type_of<int8_t>());
        auto tv = _Const(onehot + (size_t)s * n * tclasses, {n, tclasses}, NCHW, halide_type_of<int32_t>());
        auto predict = model->multioutput_forward({xv, _Scalar<int8_t>((int8_t)ascale)});
        auto info = predict[0]->getInfo();
        if (!info || info->dim.size() != 2) return -1;
        classes = info->dim[1];
        auto loss = _NITI_LOSS_SUM(predict[0], predict[1], tv, n, tclasses, 1, 1);
        auto lp = loss->readMap<float>();
        if (!lp) return -2;
        if (loss_out) loss_out[s] = lp[0];
        if (logits_out) ::memcpy(logits_out + (size_t)s * n * classes, predict[0]->readMap<int8_t>(), (size_t)n * classes);
        if (elogits_out) elogits_out[s] = predict[1]->readMap<int8_t>()[0];
        if (!sgd->step(loss)) return -3;
        if (ms_out) ms_out[s] = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    if (weights_out) {
        size_t off = 0;
        for (auto& p : model->parameters()) {
            auto pi = p->getInfo();
            ::memcpy(weights_out + off, p->readMap<int8_t>(), pi->size);
            off += pi->size;
        }
    }
    return classes;
}

} // extern "C"
