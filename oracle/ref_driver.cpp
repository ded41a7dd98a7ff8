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
// The reference reads DSP.txt / parallel.txt from the cwd with unchecked fopen
// (hexagon_implementation.cpp:33-39, CPUBackend.cpp:155-157): ref_init() writes them (2 and 0).
#include <MNN/expr/Expr.hpp>
#include <MNN/expr/ExprCreator.hpp>
#include <MNN/expr/Executor.hpp>
#include "OpGrad.hpp"
#include "MNN_generated.h"
#include <cstdio>
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
    auto dyv = _Convert(constI8(dy, {n, c, info->dim[2], info->dim[3]}), NC4HW4);
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

} // extern "C"
