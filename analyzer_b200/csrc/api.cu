// Context, error text and small queries of libtsrate.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void ts_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

int ts_cuda_fail(cudaError_t e, const char *what)
{
    ts_set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return TS_E_CUDA;
}

extern "C" const char *ts_last_error(void) { return g_err; }

int ts_reset_barrier(unsigned long long *d_ctrl, cudaStream_t stream)
{
    TS_CUDA(cudaMemsetAsync(d_ctrl, 0, 2 * sizeof(unsigned long long), stream));
    return 0;
}

extern "C" int ts_ctx_create(const ts_params *params, int device, ts_ctx **out)
{
    if (!params || !out) {
        ts_set_error("ts_ctx_create: null argument");
        return TS_E_ARG;
    }
    if (!(params->beta > 0.0) || !(params->tau >= 0.0)) {
        ts_set_error("ts_ctx_create: beta must be > 0 and tau >= 0");
        return TS_E_ARG;
    }
    int count = 0;
    TS_CUDA(cudaGetDeviceCount(&count));
    if (device < 0 || device >= count) {
        ts_set_error("ts_ctx_create: device %d of %d", device, count);
        return TS_E_ARG;
    }
    TS_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    TS_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        ts_set_error("libtsrate is built for sm_100a (B200); device %d is sm_%d%d", device, prop.major, prop.minor);
        return TS_E_UNSUPPORTED;
    }
    if (!prop.cooperativeLaunch) {
        ts_set_error("device %d cannot launch cooperative kernels", device);
        return TS_E_UNSUPPORTED;
    }
    ts_ctx *c = new ts_ctx();
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    c->params = *params;
    c->d_ctrl = nullptr;
    cudaError_t e = cudaMalloc(&c->d_ctrl, TS_CTRL_BYTES);
    if (e == cudaSuccess) e = cudaMemset(c->d_ctrl, 0, TS_CTRL_BYTES);
    if (e != cudaSuccess) {
        delete c;
        return ts_cuda_fail(e, "control words");
    }
    *out = c;
    return 0;
}

extern "C" void ts_ctx_destroy(ts_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaFree(ctx->d_ctrl);
    delete ctx;
}

extern "C" int ts_sm_count(const ts_ctx *ctx) { return ctx ? ctx->sm_count : 0; }

extern "C" int ts_error_counts(ts_ctx *ctx, int64_t *fp_errors, int64_t *value_errors, void *stream)
{
    if (!ctx) return TS_E_ARG;
    unsigned long long fp = 0, val = 0;
    TS_CUDA(cudaMemcpyAsync(&fp, ctx->d_ctrl + CTRL_ERRORS, sizeof fp, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    TS_CUDA(cudaMemcpyAsync(&val, ctx->d_ctrl + CTRL_VALUE_ERRORS, sizeof val, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    TS_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    if (fp_errors) *fp_errors = (int64_t)fp;
    if (value_errors) *value_errors = (int64_t)val;
    return 0;
}

extern "C" int64_t ts_error_count(ts_ctx *ctx, void *stream)
{
    int64_t fp = 0, val = 0;
    if (int rc = ts_error_counts(ctx, &fp, &val, stream)) return rc;
    return fp + val;
}

extern "C" int ts_reset_errors(ts_ctx *ctx, void *stream)
{
    if (!ctx) return TS_E_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    TS_CUDA(cudaMemsetAsync(ctx->d_ctrl + CTRL_ERRORS, 0, sizeof(unsigned long long), s));
    TS_CUDA(cudaMemsetAsync(ctx->d_ctrl + CTRL_STALLED, 0, sizeof(unsigned long long), s));
    TS_CUDA(cudaMemsetAsync(ctx->d_ctrl + CTRL_VALUE_ERRORS, 0, 2 * sizeof(unsigned long long), s));   // + CTRL_SHARD
    return 0;
}
