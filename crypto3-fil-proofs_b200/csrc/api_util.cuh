// Error translation at the C ABI (no exception crosses it) and the device-side
// "bad input" flag shared by the ingest kernels.
#pragma once
#include <new>
#include <string>
#include "common.cuh"

namespace b2 {

template <class Fn> int guard(Ctx* c, Fn&& fn, bool priority = false) {
    ContextGate gate(c, priority);
    try {
        cudaSetDevice(c->device);
        fn();
        return B2G16_OK;
    } catch (const ArgError& e) {
        c->last_error = e.msg;
        return e.code;
    } catch (const CudaError& e) {
        c->last_error = std::string(cudaGetErrorName(e.e)) + ": " + cudaGetErrorString(e.e) + " in " + e.what + " (" + e.file + ":" + std::to_string(e.line) + ")";
        cudaGetLastError();
        return e.e == cudaErrorMemoryAllocation ? B2G16_E_OOM : B2G16_E_CUDA;
    } catch (const std::bad_alloc&) {
        c->last_error = "host allocation failed";
        return B2G16_E_OOM;
    } catch (...) {
        c->last_error = "internal error";
        return B2G16_E_INTERNAL;
    }
}

// int flag in device memory, zeroed on the stream; check() synchronises and throws.
struct DevFlag {
    Ctx* c; DevBuf<int> buf;
    explicit DevFlag(Ctx* ctx) : c(ctx), buf(1, ctx->stream) { B2_CUDA(cudaMemsetAsync(buf.get(), 0, sizeof(int), c->stream)); }
    int* ptr() { return buf.get(); }
    void check(const char* what) {
        int v = 0;
        B2_CUDA(cudaMemcpyAsync(&v, buf.get(), sizeof(int), cudaMemcpyDeviceToHost, c->stream));
        B2_CUDA(cudaStreamSynchronize(c->stream));
        if (v & 1) throw ArgError{B2G16_E_NOT_CANONICAL, std::string(what) + ": value not canonical (>= modulus)"};
        if (v & 2) throw ArgError{B2G16_E_NOT_ON_CURVE, std::string(what) + ": point not on curve"};
        if (v & 4) throw ArgError{B2G16_E_NOT_ON_CURVE, std::string(what) + ": point not in the prime-order subgroup"};
    }
};

}  // namespace b2
