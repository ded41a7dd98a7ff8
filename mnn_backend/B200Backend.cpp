// B200Backend.cpp — Runtime / Backend plumbing of the B200 NITI plug-in (see B200Backend.hpp).
#include <map>
#include "B200Backend.hpp"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include "core/TensorUtils.hpp"

namespace MNN {
namespace B200 {

// ------------------------------------------------------------------------------------------------
ErrorCode toErrorCode(int code, const B200Backend* bn, const char* what) {
    if (code == NITI_OK) return NO_ERROR;
    const char* text = (bn && bn->ctx()) ? bn->api().last_error(bn->ctx()) : "";
    MNN_ERROR("[niti_b200] %s failed (%d): %s\n", what, code, text);
    switch (code) {
        case NITI_ERR_ARG:         return INVALID_VALUE;
        case NITI_ERR_UNSUPPORTED: return NOT_SUPPORT;
        default:                   return NO_EXECUTION;
    }
}

// ------------------------------------------------------------------------------------------------
static const B200Runtime* gLastRuntime = nullptr;      // evidence hook only (MNNNitiB200LaunchCount)
long long launchCountOfLastRuntime() {
    return gLastRuntime && gLastRuntime->ctx() ? (long long)gLastRuntime->api().launch_count(gLastRuntime->ctx()) : 0;
}

size_t B200Runtime::bucketOf(size_t bytes) {
    size_t b = 256;
    while (b < bytes) b <<= 1;                      // powers of two up to 1 MiB, then 1 MiB steps
    if (b > (1u << 20)) b = (bytes + (1u << 20) - 1) & ~(size_t)((1u << 20) - 1);
    return b;
}

B200Runtime::B200Runtime(const Backend::Info& info) : mApi(nitiApi()) {
    if (!mApi.loaded) return;
    int device = 0;
    if (const char* env = ::getenv("NITI_B200_DEVICE")) device = ::atoi(env);
    niti_ctx* ctx = nullptr;
    int rc = mApi.ctx_create(device, &ctx);
    if (rc != NITI_OK) {
        MNN_ERROR("[niti_b200] niti_ctx_create(device %d) failed (%d): no sm_100 device, and there is no CPU fallback\n",
                  device, rc);
        return;
    }
    mCtx = ctx;
    gLastRuntime = this;
}

B200Runtime::~B200Runtime() {
    if (gLastRuntime == this) gLastRuntime = nullptr;
    if (!mCtx) return;
    mApi.stream_sync(mCtx, nullptr);
    for (auto& kv : mFree) mApi.dev_free(mCtx, kv.second);
    for (auto& kv : mLive) mApi.dev_free(mCtx, kv.first);
    mApi.ctx_destroy(mCtx);
}

void* B200Runtime::alloc(size_t bytes) const {
    const size_t bucket = bucketOf(bytes);
    std::lock_guard<std::mutex> _l(mMutex);
    auto it = mFree.find(bucket);
    void* p = nullptr;
    if (it != mFree.end()) {
        p = it->second;
        mFree.erase(it);
    } else {
        if (mApi.dev_alloc(mCtx, bucket, &p) != NITI_OK) {
            // give cached blocks back to the driver and retry once
            for (auto& kv : mFree) { mApi.dev_free(mCtx, kv.second); mTotalBytes -= kv.first; }
            mFree.clear();
            if (mApi.dev_alloc(mCtx, bucket, &p) != NITI_OK) return nullptr;
        }
        mTotalBytes += bucket;
    }
    mLive[p] = bucket;
    return p;
}

void B200Runtime::free(void* ptr) const {
    std::lock_guard<std::mutex> _l(mMutex);
    auto it = mLive.find(ptr);
    if (it == mLive.end()) return;
    mFree.insert(std::make_pair(it->second, ptr));
    mLive.erase(it);
}

void B200Runtime::onGabageCollect(int level) {
    if (!mCtx || level < 100) return;               // Executor::gc(FULL) passes 100 (express/Executor.cpp:98-106)
    std::lock_guard<std::mutex> _l(mMutex);
    mApi.stream_sync(mCtx, nullptr);
    for (auto& kv : mFree) { mApi.dev_free(mCtx, kv.second); mTotalBytes -= kv.first; }
    mFree.clear();
}

float B200Runtime::onGetMemoryInMB() {
    std::lock_guard<std::mutex> _l(mMutex);
    return (float)mTotalBytes / 1024.0f / 1024.0f;
}

Backend* B200Runtime::onCreate(const BackendConfig* config) const {
    if (!mCtx) return nullptr;
    return new B200Backend(this);
}

// ------------------------------------------------------------------------------------------------
static std::map<OpType, B200Backend::Creator*>* gCreators() {
    static std::map<OpType, B200Backend::Creator*>* creators = new std::map<OpType, B200Backend::Creator*>;
    return creators;
}

bool B200Backend::addCreator(OpType t, Creator* c) {
    auto map = gCreators();
    if (map->find(t) != map->end()) {
        MNN_PRINT("Error: %d type has be added\n", t);
        return false;
    }
    map->insert(std::make_pair(t, c));
    return true;
}

B200Backend::B200Backend(const B200Runtime* runtime) : Backend(MNN_FORWARD_B200), mRuntime(runtime) {
}

B200Backend::~B200Backend() {
    for (auto& kv : mOwned) mRuntime->free(kv.first);
    for (auto& kv : mReusable) mRuntime->free(kv.second);
}

Execution* B200Backend::onCreate(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs,
                                 const MNN::Op* op) {
    auto map = gCreators();
    auto iter = map->find(op->type());
    if (iter == map->end()) {
        declinedCounts()[(int)op->type()] += 1;
        return nullptr;        // the executor retries on its CPU backup backend (express/Executor.cpp:521-529)
    }
    Execution* e = iter->second->onCreate(inputs, outputs, op, this);
    if (e != nullptr) createdCounts()[(int)op->type()] += 1;      // evidence hook: which op types really got a B200 Execution
    else declinedCounts()[(int)op->type()] += 1;
    return e;
}

std::map<int, long long>& declinedCounts() {
    static std::map<int, long long> counts;
    return counts;
}

std::map<int, long long>& createdCounts() {
    static std::map<int, long long> counts;
    return counts;
}

void B200Backend::onExecuteBegin() const {
}

void B200Backend::onExecuteEnd() const {
    // surface asynchronous kernel errors at the end of the compute, where the executor checks codes
    int rc = api().stream_sync(ctx(), stream());
    if (rc != NITI_OK) {
        MNN_ERROR("[niti_b200] stream error at onExecuteEnd: %s\n", api().last_error(ctx()));
    }
}

bool B200Backend::onAcquireBuffer(const Tensor* tensor, StorageType storageType) {
    size_t bytes = (size_t)tensor->size();
    if (bytes == 0) bytes = (size_t)tensor->getType().bytes();
    const size_t bucket = B200Runtime::bucketOf(bytes);
    void* p = nullptr;
    auto it = mReusable.find(bucket);
    if (it != mReusable.end()) {
        p = it->second;
        mReusable.erase(it);
    } else {
        p = mRuntime->alloc(bytes);
        if (!p) return false;
    }
    mOwned[p] = std::make_pair(storageType, bucket);
    auto& buffer = const_cast<Tensor*>(tensor)->buffer();
    buffer.device = (uint64_t)(uintptr_t)p;
    return true;
}

bool B200Backend::onReleaseBuffer(const Tensor* tensor, StorageType storageType) {
    if (storageType == DYNAMIC_SEPERATE) return true;       // kept until onClearBuffer (Backend.hpp:71-76)
    void* p = (void*)(uintptr_t)tensor->deviceId();
    auto it = mOwned.find(p);
    if (it == mOwned.end()) return false;
    if (storageType == STATIC) {
        mOwned.erase(it);
        mRuntime->free(p);
        return true;
    }
    // DYNAMIC: execution is stream-ordered in op order, so a block released while the plan is laid out
    // may back a later tensor of the SAME plan (MNN's dynamic-allocator rule).  It must not reach
    // another ComputeCache's backend while this plan is alive, hence the per-backend list.
    mReusable.insert(std::make_pair(it->second.second, p));
    mOwned.erase(it);
    return true;
}

bool B200Backend::onClearBuffer() {
    for (auto it = mOwned.begin(); it != mOwned.end();) {
        if (it->second.first != STATIC) {
            mRuntime->free(it->first);
            it = mOwned.erase(it);
        } else {
            ++it;
        }
    }
    for (auto& kv : mReusable) mRuntime->free(kv.second);
    mReusable.clear();
    return true;
}

static Tensor::DimensionType dimTypeOf(const Tensor* t) {
    switch (TensorUtils::getDescribe(t)->dimensionFormat) {
        case MNN_DATA_FORMAT_NHWC:   return Tensor::TENSORFLOW;
        case MNN_DATA_FORMAT_NC4HW4: return Tensor::CAFFE_C4;
        default:                     return Tensor::CAFFE;
    }
}

// Host <-> device copies.  A device tensor holds exactly the bytes the CPU backend would hold for a
// tensor of the same shape and format, so equal formats are one memcpy; different formats go through
// the reference's own host converter (MNNCPUCopyBuffer, source/core/Backend.hpp:300) on a staging tensor.
void B200Backend::onCopyBuffer(const Tensor* srcTensor, const Tensor* dstTensor) const {
    const bool srcDevice = srcTensor->deviceId() != 0 && srcTensor->host<void>() == nullptr;
    const bool dstDevice = dstTensor->deviceId() != 0 && dstTensor->host<void>() == nullptr;
    const auto srcFormat = TensorUtils::getDescribe(srcTensor)->dimensionFormat;
    const auto dstFormat = TensorUtils::getDescribe(dstTensor)->dimensionFormat;
    const bool sameFormat = srcFormat == dstFormat || srcTensor->dimensions() <= 1;
    auto& a = api();
    int rc = NITI_OK;
    if (srcDevice && dstDevice) {
        if (!sameFormat) {
            MNN_ERROR("[niti_b200] device->device copy across formats is not supported\n");
            return;
        }
        rc = a.copy_d2d(ctx(), (void*)(uintptr_t)dstTensor->deviceId(), (const void*)(uintptr_t)srcTensor->deviceId(),
                        (size_t)srcTensor->size(), stream());
    } else if (srcDevice) {
        if (sameFormat) {
            rc = a.copy_d2h(ctx(), dstTensor->host<void>(), (const void*)(uintptr_t)srcTensor->deviceId(),
                            (size_t)srcTensor->size(), stream());
            if (rc == NITI_OK) rc = a.stream_sync(ctx(), stream());
        } else {
            std::unique_ptr<Tensor> staging(new Tensor(srcTensor, dimTypeOf(srcTensor), true));
            rc = a.copy_d2h(ctx(), staging->host<void>(), (const void*)(uintptr_t)srcTensor->deviceId(),
                            (size_t)srcTensor->size(), stream());
            if (rc == NITI_OK) rc = a.stream_sync(ctx(), stream());
            if (rc == NITI_OK) MNNCPUCopyBuffer(staging.get(), dstTensor);
        }
    } else if (dstDevice) {
        if (sameFormat) {
            rc = a.copy_h2d(ctx(), (void*)(uintptr_t)dstTensor->deviceId(), srcTensor->host<void>(),
                            (size_t)dstTensor->size(), stream());
            // pageable host memory: the runtime stages the source before returning, so the caller may
            // reuse it immediately; sync anyway to keep the reference's blocking copy semantics
            if (rc == NITI_OK) rc = a.stream_sync(ctx(), stream());
        } else {
            std::unique_ptr<Tensor> staging(new Tensor(dstTensor, dimTypeOf(dstTensor), true));
            MNNCPUCopyBuffer(srcTensor, staging.get());
            rc = a.copy_h2d(ctx(), (void*)(uintptr_t)dstTensor->deviceId(), staging->host<void>(),
                            (size_t)dstTensor->size(), stream());
            if (rc == NITI_OK) rc = a.stream_sync(ctx(), stream());
        }
    } else {
        MNNCPUCopyBuffer(srcTensor, dstTensor);
    }
    if (rc != NITI_OK) {
        MNN_ERROR("[niti_b200] onCopyBuffer failed (%d): %s\n", rc, a.last_error(ctx()));
    }
}

// ------------------------------------------------------------------------------------------------
Runtime* B200RuntimeCreator::onCreate(const Backend::Info& info) const {
    auto rt = new B200Runtime(info);
    if (!rt->valid()) {
        delete rt;
        return nullptr;
    }
    return rt;
}

bool B200RuntimeCreator::onValid(Backend::Info& info) const {
    info.mode = Backend::Info::DIRECT;
    return nitiApi().loaded;
}

} // namespace B200
} // namespace MNN
