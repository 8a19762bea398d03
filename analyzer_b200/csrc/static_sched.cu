// The scheduler of sched.cu for callers that must not allocate device memory.
#include <string.h>

#include <mutex>

#include "common.cuh"

// ---------------------------------------------------------------------------
// ts_schedule_static: the same scheduler for callers that must not allocate.
//
// Who needs it: a rating worker that runs inside someone else's process -- a
// plugin in a service whose device memory is handed out by the host
// application's own pool, a job under a fixed memory budget, a component
// running next to a framework that has reserved all of HBM -- cannot call
// cudaMalloc (or torch's allocator) for the scheduler's workspace, and
// ts_ctx_create itself allocates its control words.  Such a caller still owns
// its input and output buffers; only the scratch is the problem.  Here the
// scratch (sort ping-pong buffers, per-block histograms, successor links,
// predecessor counts, the frontier queue and the barrier/tail control words)
// lives in one file-scope __device__ array, which every GPU that loads
// libtsrate.so holds from load time until the process exits.
//
// Limits and rules:
//   * n_matches * slots <= TS_STATIC_MAX_SLOT_ENTRIES (65 536: more than ten of
//     the reference worker's 500-match batches, worker.py:18); larger batches
//     are rejected with TS_E_WORKSPACE -- use ts_schedule with your own workspace.
//   * The first call on a device looks up the array's device address
//     (cudaGetSymbolAddress, with that device current) and the SM count, and
//     caches both; make that first call before any stream capture.
//   * Every call shares the one array and leaves its contents behind; whatever
//     must start at zero (predecessor counts, rounds, barrier and queue words) is
//     cleared by the call itself, on the caller's stream.  Two calls whose GPU
//     work overlaps (different streams, different threads) race on the array:
//     serialise them.
//   * No cudaMalloc / cudaFree / allocating library call is made.
// ---------------------------------------------------------------------------
namespace {

constexpr size_t kStaticCtrlBytes = (TS_CTRL_BYTES + 255) & ~(size_t)255;   // control words first, then the workspace
constexpr size_t kStaticScratchBytes = 2u << 20;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

struct StaticDevice {
    bool ready;
    unsigned char *scratch;   // device address of g_static_scratch on this device
    int sm_count;
};
StaticDevice g_static_dev[64];
std::mutex g_static_lock;      // first calls from two host threads fill the cache one after the other

// once per device, with that device current
int static_setup(int device, StaticDevice **out)
{
    if (device < 0 || device >= 64) {
        ts_set_error("ts_schedule_static: device %d out of range", device);
        return TS_E_ARG;
    }
    StaticDevice &d = g_static_dev[device];
    std::lock_guard<std::mutex> guard(g_static_lock);
    if (!d.ready) {
        TS_CUDA(cudaSetDevice(device));
        void *ptr = nullptr;
        // in host code `g_static_scratch` is only a handle for the symbol, not a
        // device address: ask the runtime for the address on the current device
        TS_CUDA(cudaGetSymbolAddress(&ptr, g_static_scratch));
        int sms = 0, coop = 0, major = 0;
        TS_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        TS_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device));
        TS_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
        if (major != 10 || !coop) {
            ts_set_error("ts_schedule_static: device %d is not a B200-class (sm_100) device", device);
            return TS_E_UNSUPPORTED;
        }
        d.scratch = (unsigned char *)ptr;
        d.sm_count = sms;
        d.ready = true;
    }
    *out = &d;
    return 0;
}

}  // namespace

extern "C" int64_t ts_schedule_static_max_slot_entries(void) { return TS_STATIC_MAX_SLOT_ENTRIES; }

extern "C" int ts_schedule_static(int device, const uint32_t *ids, const uint32_t *meta, int64_t n_matches,
                                  int32_t slots, int64_t n_players, const int8_t mode_col[16],
                                  uint32_t *round, uint32_t *order, uint32_t *round_offsets,
                                  int64_t offsets_cap, uint32_t *ids_sorted, uint32_t *meta_sorted,
                                  ts_sched_info *info, uint32_t *n_rounds_host, void *stream)
{
    if (n_matches < 0 || slots < 2 || slots > TS_MAX_SLOTS || (slots & 1)) {
        ts_set_error("ts_schedule_static: bad n_matches / slots");
        return TS_E_ARG;
    }
    if (n_matches * (int64_t)slots > TS_STATIC_MAX_SLOT_ENTRIES) {
        ts_set_error("ts_schedule_static: n_matches * slots = %lld exceeds the built-in scratch (%d slot entries); "
                     "use ts_schedule with a caller-provided workspace",
                     (long long)(n_matches * (int64_t)slots), TS_STATIC_MAX_SLOT_ENTRIES);
        return TS_E_WORKSPACE;
    }
    StaticDevice *d = nullptr;
    int rc = static_setup(device, &d);
    if (rc) return rc;
    const int64_t need = ts_schedule_workspace_bytes(n_matches, slots);
    if (need + (int64_t)kStaticCtrlBytes > (int64_t)kStaticScratchBytes) {
        ts_set_error("ts_schedule_static: internal scratch too small (%lld bytes needed)", (long long)need);
        return TS_E_WORKSPACE;
    }
    // a context on the stack: the control words are the head of the static array
    ts_ctx ctx;
    ctx.device = device;
    ctx.sm_count = d->sm_count;
    memset(&ctx.params, 0, sizeof ctx.params);
    ctx.d_ctrl = (unsigned long long *)d->scratch;
    return ts_schedule(&ctx, ids, meta, n_matches, slots, n_players, mode_col, round, order, round_offsets,
                       offsets_cap, ids_sorted, meta_sorted, info, d->scratch + kStaticCtrlBytes,
                       (int64_t)(kStaticScratchBytes - kStaticCtrlBytes), n_rounds_host, stream);
}
