// B200Backend.hpp — MNN RuntimeCreator / Runtime / Backend for the B200 NITI int8 path.
//
// Implements the reference's backend contract (execution-engine/source/core/Backend.hpp:113-298):
//   RuntimeCreator::onCreate/onValid  -> B200RuntimeCreator
//   Runtime::onCreate/onGabageCollect/onGetCompilerType/onGetMemoryInMB -> B200Runtime
//   Backend::onCreate/onAcquireBuffer/onReleaseBuffer/onClearBuffer/onCopyBuffer/onExecuteBegin/End -> B200Backend
// and the per-op creator pattern of the CPU backend (source/backend/cpu/CPUBackend.hpp:85-91,179-183).
// The executor makes a fresh Backend per ComputeCache (express/Executor.cpp:657-668), so the CUDA
// context and the device-memory pool live in the Runtime.
//
// Device tensors keep the byte layout the CPU backend gives the same tensor (int8 "NC4HW4" =
// [c/4][n][h*w][4], NCHW, NHWC), so copies across the backend boundary are plain memcpys and the
// Executions convert to the kernels' NHWC/OHWI operands on the device.
#ifndef MNN_B200_BACKEND_HPP
#define MNN_B200_BACKEND_HPP

#include <map>
#include <mutex>
#include <vector>
#include "core/Backend.hpp"
#include "core/Execution.hpp"
#include "NitiB200Api.hpp"

namespace MNN {
namespace B200 {

#ifndef MNN_FORWARD_B200
#define MNN_FORWARD_B200 MNN_FORWARD_USER_0      // free plug-in slot, include/MNN/MNNForwardType.h:35-39
#endif

class B200Runtime : public Runtime {
public:
    explicit B200Runtime(const Backend::Info& info);
    virtual ~B200Runtime();
    bool valid() const { return mCtx != nullptr; }

    virtual Backend* onCreate(const BackendConfig* config = nullptr) const override;
    virtual void onGabageCollect(int level) override;
    virtual CompilerType onGetCompilerType() const override { return Compiler_Loop; }   // same as CPU/CUDA
    virtual float onGetMemoryInMB() override;

    niti_ctx* ctx() const { return mCtx; }
    const NitiB200Api& api() const { return mApi; }
    // size-bucketed caching allocator over niti_dev_alloc (cudaMalloc is too slow for the executor's
    // per-step graph rebuild, express/Executor.cpp:371-611)
    void* alloc(size_t bytes) const;
    void  free(void* ptr) const;
    static size_t bucketOf(size_t bytes);

private:
    const NitiB200Api& mApi;
    niti_ctx* mCtx = nullptr;
    mutable std::mutex mMutex;
    mutable std::multimap<size_t, void*> mFree;      // bucket size -> pointer
    mutable std::map<void*, size_t> mLive;           // pointer -> bucket size
    mutable size_t mTotalBytes = 0;
};

class B200Backend : public Backend {
public:
    explicit B200Backend(const B200Runtime* runtime);
    virtual ~B200Backend();

    class Creator {
    public:
        virtual ~Creator() = default;
        virtual Execution* onCreate(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs,
                                    const MNN::Op* op, Backend* backend) const = 0;
    };
    static bool addCreator(OpType t, Creator* c);

    virtual Execution* onCreate(const std::vector<Tensor*>& inputs, const std::vector<Tensor*>& outputs,
                                const MNN::Op* op) override;
    virtual void onExecuteBegin() const override;
    virtual void onExecuteEnd() const override;
    virtual bool onAcquireBuffer(const Tensor* tensor, StorageType storageType) override;
    virtual bool onReleaseBuffer(const Tensor* tensor, StorageType storageType) override;
    virtual bool onClearBuffer() override;
    virtual void onCopyBuffer(const Tensor* srcTensor, const Tensor* dstTensor) const override;

    const B200Runtime* runtime() const { return mRuntime; }
    niti_ctx* ctx() const { return mRuntime->ctx(); }
    const NitiB200Api& api() const { return mRuntime->api(); }
    niti_stream stream() const { return nullptr; }       // the executor is single-threaded and serial (Executor.cpp:767-796)
    // How many Executions of this backend asked with the same key before (0 for the first).  The executor builds a fresh
    // Backend for every ComputeCache and creates its Executions in graph order, so "the k-th op of this geometry" names the
    // same layer in every training step: repeated blocks (VGG's two 512->512 layers, a ResNet stage) get predictor slots of
    // their own instead of sharing one.
    int occurrenceOf(unsigned long long key) { return mOccurrence[key]++; }

private:
    const B200Runtime* mRuntime;
    std::map<void*, std::pair<StorageType, size_t>> mOwned;   // live buffers of this backend: type, bucket size
    std::multimap<size_t, void*> mReusable;                    // DYNAMIC blocks released during planning
    std::map<unsigned long long, int> mOccurrence;
};

class B200RuntimeCreator : public RuntimeCreator {
public:
    virtual Runtime* onCreate(const Backend::Info& info) const override;
    virtual bool onValid(Backend::Info& info) const override;
};

template <class T>
class B200CreatorRegister {
public:
    explicit B200CreatorRegister(OpType type) { B200Backend::addCreator(type, new T); }
};
#define REGISTER_B200_OP_CREATOR(name, opType) \
    static ::MNN::B200::B200CreatorRegister<name> _b200_register_##name##_##opType(opType)

long long launchCountOfLastRuntime();
std::map<int, long long>& createdCounts();
// op types the backend declined (no creator, or the creator returned nullptr): they ran on the executor's CPU backup backend
std::map<int, long long>& declinedCounts();   // Executions created so far, by op type

// NITI_* -> ErrorCode
ErrorCode toErrorCode(int nitiCode, const B200Backend* bn, const char* what);

} // namespace B200
} // namespace MNN
#endif
