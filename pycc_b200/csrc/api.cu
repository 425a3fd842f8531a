// Context management of libpyccb200.so (see include/pyccb200.h).
#include "common.cuh"

thread_local std::string g_create_error;

extern "C" int pccb_abi_version(void) { return PCCB_ABI_VERSION; }

extern "C" int pccb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

extern "C" int pccb_create(int device, pccb_ctx** out) {
    if (!out) return PCCB_ERR_ARG;
    *out = nullptr;
    int n = pccb_device_count();
    if (n <= 0) {
        g_create_error = "no CUDA device: libpyccb200 has no CPU fallback";
        return PCCB_ERR_NODEV;
    }
    if (device < 0 || device >= n) {
        g_create_error = "device index out of range";
        return PCCB_ERR_ARG;
    }
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) {
        g_create_error = std::string("cudaSetDevice failed: ") + cudaGetErrorString(e);
        return PCCB_ERR_CUDA;
    }
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) {
        g_create_error = std::string("cudaGetDeviceProperties failed: ") + cudaGetErrorString(e);
        return PCCB_ERR_CUDA;
    }
    if (prop.major != 10) {
        g_create_error = "device is not sm_100 (Blackwell B200); this library ships sm_100a code only";
        return PCCB_ERR_NODEV;
    }
    pccb_ctx* ctx = new pccb_ctx();
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    e = cudaMallocHost(&ctx->result_host, sizeof(double) * 64);
    if (e != cudaSuccess) {
        g_create_error = std::string("cudaMallocHost failed: ") + cudaGetErrorString(e);
        delete ctx;
        return PCCB_ERR_CUDA;
    }
    *out = ctx;
    return PCCB_OK;
}

extern "C" int pccb_destroy(pccb_ctx* ctx) {
    if (!ctx) return PCCB_ERR_ARG;
    if (ctx->partial_dev) cudaFree(ctx->partial_dev);
    if (ctx->result_host) cudaFreeHost(ctx->result_host);
    delete ctx;
    return PCCB_OK;
}

extern "C" const char* pccb_last_error(pccb_ctx* ctx) {
    if (!ctx) return g_create_error.c_str();
    return ctx->err.c_str();
}

extern "C" int pccb_set_stream(pccb_ctx* ctx, void* cuda_stream) {
    if (!ctx) return PCCB_ERR_ARG;
    ctx->stream = static_cast<cudaStream_t>(cuda_stream);
    return PCCB_OK;
}

extern "C" int pccb_set_workspace(pccb_ctx* ctx, void* ws_dev, size_t bytes) {
    if (!ctx) return PCCB_ERR_ARG;
    ctx->ws = ws_dev;
    ctx->ws_bytes = ws_dev ? bytes : 0;
    return PCCB_OK;
}

extern "C" int pccb_set_contract_impl(pccb_ctx* ctx, int impl) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, impl == PCCB_IMPL_TENSOR || impl == PCCB_IMPL_NAIVE, "unknown contraction impl %d", impl);
    ctx->impl = impl;
    return PCCB_OK;
}

extern "C" int pccb_launch_count(pccb_ctx* ctx, int64_t* n_out) {
    if (!ctx || !n_out) return PCCB_ERR_ARG;
    *n_out = ctx->launches;
    return PCCB_OK;
}

extern "C" int pccb_synchronize(pccb_ctx* ctx) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PCCB_OK;
}
