// fmpnp_capi.cu -- the C ABI declared in include/fmpnp.h: context, workspace, argument checks,
// launch geometry.  No compute happens on the host; without a CUDA device fmpnp_create() fails.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "fmpnp_kernels.cuh"

using namespace fmpnp;

struct Geometry_last {
    int grid, cluster, ring_bytes, ctas;
};

struct fmpnp_ctx {
    int device = 0;
    int sm_count = 0;
    int max_smem_optin = 0;
    char err[512] = {0};
    // device copy of the problem descriptors + pinned staging
    QueryDev* d_queries = nullptr;
    QueryDev* h_queries[2] = {nullptr, nullptr};  // pinned staging, alternating between consecutive calls
    size_t cap_queries = 0;
    cudaEvent_t staged[2] = {nullptr, nullptr};
    bool staged_pending[2] = {false, false};
    int stage_slot = 0;
    // per-CTA scratch
    double* d_rec = nullptr;
    uint8_t* d_supp = nullptr;
    uint8_t* d_inl = nullptr;   // inlier masks of the fused find_inliers pass
    size_t cap_points = 0;
    unsigned int* d_counter = nullptr;
    // geometry (overridable through the environment for tuning)
    int ring_bytes = 0;   // 0 = automatic (see geometry())
    int depth = 0;
    int ctas_per_sm = 0;
    int cluster = 0;
    int scheduler = 1;   // 1: task queue (default), 0: one CTA / cluster per query
    int watchdog_s = 20;   // FMPNP_WATCHDOG_S: seconds a CTA may wait for a task before the launch is abandoned (0 = off)
    int small_batch_clusters = 0;  // 1: a handful of queries (B <= 4) go to clusters even with the scheduler on
    int units = 0;       // tasks per evaluation (0 = automatic)
    QueryState* d_states = nullptr;
    double* d_part = nullptr;
    size_t cap_states = 0;
    unsigned long long* d_queue = nullptr;
    uint32_t cap_queue = 0;
    unsigned int* d_sched = nullptr;  // head, tail, done, error
    int smem_configured = -1;
    Geometry_last last = {0, 0, 0, 0};
    int64_t launches = 0;
    // record reuse: generations handed out so far (tags of `d_rec` hold generation << 32 | cell, never generation 0)
    uint64_t reuse_next = 1;
    int reuse = 1;  // FMPNP_REUSE=0: gather every row at every evaluation, whatever the options say
    int light_miss_den = 64;  // FMPNP_LIGHT_MISS_DEN: the light route is left while an evaluation gathers more than 1 / this of its points
    int l2_policy = 1;  // FMPNP_L2_POLICY=0: no L2 eviction preference for gathered rows / staged records
    int light = -1;  // FMPNP_LIGHT: 1 / 0 = the light evaluations of a launch with record reuse always / never take their own route (fmpnp_light.cuh); -1 = automatic
    // One workspace per context: a call issued on another stream than the previous one waits for that call's
    // kernels first (cudaStreamWaitEvent), so that two launches of a context never overlap on the device.
    // fmpnp_options.keep_records: what the previous refinement launch worked on (valid while nothing else has touched
    // the workspace since)
    bool prev_valid = false;
    int32_t prev_B = 0, prev_n_max = 0;
    uint32_t prev_base = 0;
    std::vector<QueryDev> prev_queries;
    cudaEvent_t last_done = nullptr;
    cudaStream_t last_stream = nullptr;
    bool last_pending = false;
};

namespace {

constexpr int kNoallocRingBytes = 96 * 1024;  // staging ring of fmpnp_compute_cost_batch_noalloc's CTAs (a multiple of 8 x 128)

int fail(fmpnp_ctx* ctx, int code, const char* fmt, ...) {
    if (ctx) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(ctx->err, sizeof(ctx->err), fmt, ap);
        va_end(ap);
    }
    return code;
}

#define FMPNP_CUDA(ctx, call)                                                                           \
    do {                                                                                                \
        cudaError_t e__ = (call);                                                                       \
        if (e__ != cudaSuccess)                                                                         \
            return fail((ctx), e__ == cudaErrorMemoryAllocation ? FMPNP_ERR_OUT_OF_MEMORY : FMPNP_ERR_CUDA, \
                        "%s failed: %s", #call, cudaGetErrorString(e__));                               \
    } while (0)

int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    if (!v || !*v) return dflt;
    return atoi(v);
}

int validate_problem(fmpnp_ctx* ctx, const fmpnp_problem& p, int b, bool need_grad) {
    if (p.N < 0 || p.C < 1 || p.Hf < 1 || p.Wf < 1 || p.im_width < 1 || p.im_height < 1)
        return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "problem %d: bad sizes N=%d C=%d Hf=%d Wf=%d im=%dx%d", b, p.N, p.C,
                    p.Hf, p.Wf, p.im_width, p.im_height);
    if (p.Hf > 32767 || p.Wf > 32767)
        return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "problem %d: Hf, Wf must be <= 32767 (int16 cell index of the reference)", b);
    if (!p.F || (p.N > 0 && (!p.ref || !p.pts3D)))
        return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "problem %d: null F / ref / pts3D", b);
    if (need_grad && (!p.Gx || !p.Gy))
        return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "problem %d: null Gx / Gy", b);
    return FMPNP_SUCCESS;
}

// The workspace (records, flags, states, queue, counters) belongs to the context, not to a stream: a call on a
// different stream than the previous call's is ordered behind it on the device.  Same stream: already ordered.
int order_after_previous(fmpnp_ctx* ctx, cudaStream_t stream) {
    if (ctx->last_pending && stream != ctx->last_stream) FMPNP_CUDA(ctx, cudaStreamWaitEvent(stream, ctx->last_done, 0));
    return FMPNP_SUCCESS;
}

int mark_launched(fmpnp_ctx* ctx, cudaStream_t stream) {
    FMPNP_CUDA(ctx, cudaEventRecord(ctx->last_done, stream));
    ctx->last_stream = stream;
    ctx->last_pending = true;
    return FMPNP_SUCCESS;
}

// copies the descriptors to the device (through pinned staging) and sizes the per-CTA scratch
int stage(fmpnp_ctx* ctx, const fmpnp_problem* probs, int32_t B, int grid, int32_t* n_max_out, cudaStream_t stream) {
    if ((size_t)B > ctx->cap_queries) {
        for (int i = 0; i < 2; ++i) {
            if (ctx->staged_pending[i]) {
                FMPNP_CUDA(ctx, cudaEventSynchronize(ctx->staged[i]));
                ctx->staged_pending[i] = false;
            }
        }
        size_t cap = ctx->cap_queries ? ctx->cap_queries : 256;
        while (cap < (size_t)B) cap *= 2;
        if (ctx->d_queries) FMPNP_CUDA(ctx, cudaFree(ctx->d_queries));
        ctx->d_queries = nullptr;
        ctx->cap_queries = 0;
        for (int i = 0; i < 2; ++i) {
            if (ctx->h_queries[i]) FMPNP_CUDA(ctx, cudaFreeHost(ctx->h_queries[i]));
            ctx->h_queries[i] = nullptr;
        }
        FMPNP_CUDA(ctx, cudaMalloc(&ctx->d_queries, cap * sizeof(QueryDev)));
        for (int i = 0; i < 2; ++i) FMPNP_CUDA(ctx, cudaMallocHost(&ctx->h_queries[i], cap * sizeof(QueryDev)));
        ctx->cap_queries = cap;
    }
    const int slot = ctx->stage_slot;
    ctx->stage_slot ^= 1;
    if (ctx->staged_pending[slot]) {  // the H2D copy of the call before the previous one must have left this buffer
        FMPNP_CUDA(ctx, cudaEventSynchronize(ctx->staged[slot]));
        ctx->staged_pending[slot] = false;
    }
    int32_t n_max = 1;
    for (int32_t b = 0; b < B; ++b) {
        const fmpnp_problem& p = probs[b];
        QueryDev& q = ctx->h_queries[slot][b];
        q.F = p.F;
        q.Gx = p.Gx;
        q.Gy = p.Gy;
        q.ref = p.ref;
        q.pts = p.pts3D;
        q.active = p.active;
        memcpy(q.K, p.K, sizeof(q.K));
        memcpy(q.R0, p.R0, sizeof(q.R0));
        memcpy(q.t0, p.t0, sizeof(q.t0));
        q.N = p.N;
        q.C = p.C;
        q.Hf = p.Hf;
        q.Wf = p.Wf;
        q.im_w = p.im_width;
        q.im_h = p.im_height;
        if (p.N > n_max) n_max = p.N;
    }
    FMPNP_CUDA(ctx, cudaMemcpyAsync(ctx->d_queries, ctx->h_queries[slot], (size_t)B * sizeof(QueryDev),
                                    cudaMemcpyHostToDevice, stream));
    FMPNP_CUDA(ctx, cudaEventRecord(ctx->staged[slot], stream));
    ctx->staged_pending[slot] = true;
    const size_t need = (size_t)grid * (size_t)n_max;
    if (need > ctx->cap_points) {
        if (ctx->d_rec) FMPNP_CUDA(ctx, cudaFree(ctx->d_rec));
        if (ctx->d_supp) FMPNP_CUDA(ctx, cudaFree(ctx->d_supp));
        if (ctx->d_inl) FMPNP_CUDA(ctx, cudaFree(ctx->d_inl));
        ctx->d_rec = nullptr;
        ctx->d_supp = nullptr;
        ctx->d_inl = nullptr;
        ctx->cap_points = 0;
        ctx->prev_valid = false;
        FMPNP_CUDA(ctx, cudaMalloc(&ctx->d_rec, need * kRecDoubles * sizeof(double)));
        FMPNP_CUDA(ctx, cudaMalloc(&ctx->d_supp, need));
        FMPNP_CUDA(ctx, cudaMalloc(&ctx->d_inl, need));
        // no record tag of the new buffer may look like a generation that is handed out later
        FMPNP_CUDA(ctx, cudaMemsetAsync(ctx->d_rec, 0, need * kRecDoubles * sizeof(double), stream));
        ctx->cap_points = need;
    }
    *n_max_out = n_max;
    return FMPNP_SUCCESS;
}

// Launch geometry.
//   * Default: the task scheduler.  An evaluation is cut into U tasks (1..128) so that about three tasks per
//     CTA slot exist (tasks keep at least 64 points, 32 when the launch would not fill the SMs otherwise); 2 CTAs per SM with a 96 KB staging ring each, or one CTA per SM with the whole shared
//     memory when there is at most one task per SM.
//   * An explicit cluster_size or FMPNP_SCHEDULER=0: one CTA, or a thread-block cluster of 2..16 CTAs, per
//     query (the points are split, the normal equations combined through distributed shared memory).
// FMPNP_RING_KB / FMPNP_MIN_DEPTH / FMPNP_CTAS_PER_SM / FMPNP_CLUSTER / FMPNP_UNITS override the choice
// (tuning and tests only).
struct Geometry {
    int grid, ring_bytes, ring_per_warp, min_depth, cluster, units;
};

// Light evaluations (fmpnp_light.cuh) are run as a whole by one CTA.  Worth it when there are enough queries to keep the
// CTAs busy that way AND the evaluation has a ratio test (cut into tasks it needs two rounds through the queue: the
// limit is a maximum over all points); without one, tasks keep every CTA slot busy and measure faster
// (profiles/r02/light_route.md).  Same bits either way.
bool light_route(const fmpnp_ctx* ctx, int32_t B, bool ratio) {
    return ctx->light >= 0 ? ctx->light != 0 : (ratio && (int64_t)B * 2 >= ctx->sm_count);
}

int32_t max_points(const fmpnp_problem* probs, int32_t B) {
    int32_t n_max = 1;
    for (int32_t b = 0; b < B; ++b)
        if (probs[b].N > n_max) n_max = probs[b].N;
    return n_max;
}

int geometry(fmpnp_ctx* ctx, int32_t B, int32_t n_max, int requested_cluster, Geometry* g, bool reuse = false, bool ratio = false) {
    // the light evaluations of a launch with record reuse take their own route (LaunchParams::light): the tasks are then
    // sized for the evaluations that gather
    if (reuse && light_route(ctx, B, ratio)) reuse = false;
    int ring = ctx->ring_bytes, d = ctx->depth, per_sm = ctx->ctas_per_sm;
    int S = requested_cluster > 0 ? requested_cluster : ctx->cluster;
    // Task scheduler (default): an evaluation is cut into U tasks so that about three tasks per CTA slot are
    // available (load balance) while a task still gives every warp a whole batch; no clusters needed.
    // An explicit cluster_size (options / FMPNP_CLUSTER) or FMPNP_SCHEDULER=0 selects the per-query path.
    // (FMPNP_SMALL_BATCH_CLUSTERS=1 sends B <= 4 to clusters as well: that was faster while an evaluation could
    // only be cut into 16 tasks; with 128 the scheduler refines one cfg3 query in 0.73 ms against 1.99 ms.)
    bool tasks = ctx->scheduler == 1 && S == 0;
    if (tasks && ctx->units <= 0 && ctx->small_batch_clusters && (int64_t)B * 32 <= ctx->sm_count) tasks = false;
    int U = 1;
    if (tasks) {
        U = ctx->units;
        if (U <= 0) {
            U = 1;  // ... with at least 64 points per task, or 32 when the tasks would not even fill the CTA slots
            // (with record reuse most evaluations gather next to nothing and pay per task: half as many, larger tasks)
            const int64_t per_slot_x2 = reuse ? 3 : 6;  // tasks wanted per CTA slot, times two
            while (U < kRanges && (int64_t)B * U * 2 < per_slot_x2 * 2 * (int64_t)ctx->sm_count && (int64_t)n_max >= 128 * (int64_t)U) U *= 2;
            while (U < kRanges && (int64_t)B * U < 2 * (int64_t)ctx->sm_count && (int64_t)n_max >= 64 * (int64_t)U) U *= 2;
        }
        if (U < 1 || U > kRanges || (U & (U - 1)) != 0)
            return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "FMPNP_UNITS must be a power of two between 1 and %d; got %d", kRanges, U);
        S = 1;
    }
    g->units = tasks ? U : 0;
    const int64_t work = (int64_t)B * U;  // CTAs that can be busy at once
    const int static_b = (int)refine_static_smem_bytes();
    const int budget = ctx->max_smem_optin - static_b - 1024;
    if (S != 0 && S != 1 && S != 2 && S != 4 && S != 8 && S != 16)
        return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "cluster_size must be 0 (auto), 1, 2, 4, 8 or 16; got %d", S);
    const bool auto_cluster = S == 0;
    if (auto_cluster && !tasks) {
        S = 1;
        while (S < 16 && (int64_t)B * S * 2 <= 2 * (int64_t)ctx->sm_count) S *= 2;
    }
    if (d <= 0) d = 2;
    if (d > kMaxDepth) d = kMaxDepth;
    for (;;) {
        int ps = per_sm > 0 ? (per_sm > 2 ? 2 : per_sm) : ((work * S <= ctx->sm_count) ? 1 : 2);
        int r = ring > 0 ? ring : (ps == 1 ? budget : 96 * 1024);
        if (r > budget) r = budget;
        if (r < kMinRingPerWarp * kWarps) r = kMinRingPerWarp * kWarps;  // the reductions use a warp's slice as scratch
        const int per_warp = (r / kWarps) & ~127;
        if (per_warp < 2 * 128) return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "staging ring too small (%d B)", r);
        r = per_warp * kWarps;
        // The kernels' function attributes belong to the process (per device), not to a context: every context
        // sets the same values -- the largest ring any launch may ask for, non-portable cluster sizes allowed --
        // so that one context never invalidates what another one relies on.
        if (ctx->smem_configured < 0) {
            FMPNP_CUDA(ctx, refine_set_max_dynamic_smem(((budget / kWarps) & ~127) * kWarps));
            FMPNP_CUDA(ctx, refine_allow_nonportable_clusters(1));
            ctx->smem_configured = 1;
        }
        int clusters = ctx->sm_count * ps / S;
        if (S > 1) {
            const int fit = max_active_clusters(S, r);
            if (fit < 1) {  // such a cluster cannot be scheduled: halve it
                if (!auto_cluster && S == requested_cluster)
                    return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "a cluster of %d CTAs with %d B of shared memory cannot be scheduled", S, r);
                S /= 2;
                continue;
            }
            if (clusters > fit) clusters = fit;
        }
        if (clusters > work) clusters = (int)work;
        if (clusters < 1) clusters = 1;
        g->grid = clusters * S;
        g->ring_bytes = r;
        g->ring_per_warp = per_warp;
        g->min_depth = d;
        g->cluster = S;
        return FMPNP_SUCCESS;
    }
}

void fill_common(LaunchParams& lp, fmpnp_ctx* ctx, int32_t B, int32_t n_max, const Geometry& g) {
    memset(&lp, 0, sizeof(lp));
    lp.queries = ctx->d_queries;
    lp.B = B;
    lp.rec = ctx->d_rec;
    lp.supp = ctx->d_supp;
    lp.inl = ctx->d_inl;
    lp.n_max = n_max;
    lp.work_counter = ctx->d_counter;
    lp.ring_per_warp = g.ring_per_warp;
    lp.min_depth = g.min_depth;
    lp.cluster_size = g.cluster;
    lp.scheduler = g.units > 0 ? 1 : 0;
    lp.l2_policy = ctx->l2_policy;
    lp.light_miss_den = ctx->light_miss_den;
}

// Task scheduler workspace: per-query LM states, the task FIFO and its counters; reset on the stream.
int prepare_tasks(fmpnp_ctx* ctx, int32_t B, const Geometry& g, LaunchParams& lp, cudaStream_t stream) {
    if (g.units <= 0) return FMPNP_SUCCESS;
    if ((size_t)B > ctx->cap_states) {
        if (ctx->d_states) FMPNP_CUDA(ctx, cudaFree(ctx->d_states));
        ctx->d_states = nullptr;
        ctx->cap_states = 0;
        if (ctx->d_part) FMPNP_CUDA(ctx, cudaFree(ctx->d_part));
        ctx->d_part = nullptr;
        FMPNP_CUDA(ctx, cudaMalloc(&ctx->d_states, (size_t)B * sizeof(QueryState)));
        FMPNP_CUDA(ctx, cudaMalloc(&ctx->d_part, (size_t)B * kRanges * 32 * sizeof(double)));  // 32 KB per query
        ctx->cap_states = (size_t)B;
    }
    uint64_t want = 2ull * (uint64_t)B * (uint64_t)g.units;  // at most B * units tasks are ever queued at once
    uint32_t cap = 1024;
    while (cap < want) cap <<= 1;
    if (cap > ctx->cap_queue) {
        if (ctx->d_queue) FMPNP_CUDA(ctx, cudaFree(ctx->d_queue));
        ctx->d_queue = nullptr;
        ctx->cap_queue = 0;
        FMPNP_CUDA(ctx, cudaMalloc(&ctx->d_queue, (size_t)cap * sizeof(unsigned long long)));
        ctx->cap_queue = cap;
    }
    if (!ctx->d_sched) FMPNP_CUDA(ctx, cudaMalloc(&ctx->d_sched, 8 * sizeof(unsigned int)));
    FMPNP_CUDA(ctx, cudaMemsetAsync(ctx->d_queue, 0xff, (size_t)ctx->cap_queue * sizeof(unsigned long long), stream));
    FMPNP_CUDA(ctx, cudaMemsetAsync(ctx->d_sched, 0, 8 * sizeof(unsigned int), stream));
    lp.sched.queue = ctx->d_queue;
    lp.sched.head = ctx->d_sched;
    lp.sched.tail = ctx->d_sched + 1;
    lp.sched.done = ctx->d_sched + 2;
    lp.sched.error = ctx->d_sched + 3;
    lp.sched.started = ctx->d_sched + 4;
    lp.sched.wait_limit_ns = (unsigned long long)ctx->watchdog_s * 1000000000ull;
    lp.sched.cap = ctx->cap_queue;
    lp.sched.units = g.units;
    lp.sched.states = ctx->d_states;
    lp.sched.part = ctx->d_part;
    return FMPNP_SUCCESS;
}

}  // namespace

extern "C" {

int fmpnp_version(void) { return FMPNP_VERSION; }

int fmpnp_struct_size(int which) {
    switch (which) {
        case 0: return (int)sizeof(fmpnp_problem);
        case 1: return (int)sizeof(fmpnp_options);
        case 2: return (int)sizeof(fmpnp_result);
        case 3: return (int)sizeof(fmpnp_history);
        default: return -1;
    }
}

int fmpnp_create(fmpnp_ctx** out, int device) {
    if (!out) return FMPNP_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) return FMPNP_ERR_NO_DEVICE;
    if (device < 0 || device >= count) return FMPNP_ERR_INVALID_ARGUMENT;
    fmpnp_ctx* ctx = new (std::nothrow) fmpnp_ctx();
    if (!ctx) return FMPNP_ERR_OUT_OF_MEMORY;
    ctx->device = device;
    if (cudaSetDevice(device) != cudaSuccess ||
        cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device) != cudaSuccess ||
        cudaDeviceGetAttribute(&ctx->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device) != cudaSuccess ||
        cudaMalloc(&ctx->d_counter, sizeof(unsigned int)) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->staged[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->staged[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->last_done, cudaEventDisableTiming) != cudaSuccess) {
        delete ctx;
        return FMPNP_ERR_CUDA;
    }
    ctx->ring_bytes = env_int("FMPNP_RING_KB", 0) * 1024;
    ctx->depth = env_int("FMPNP_MIN_DEPTH", 0);
    ctx->cluster = env_int("FMPNP_CLUSTER", 0);
    ctx->scheduler = env_int("FMPNP_SCHEDULER", 1);
    ctx->units = env_int("FMPNP_UNITS", 0);
    ctx->small_batch_clusters = env_int("FMPNP_SMALL_BATCH_CLUSTERS", 0);
    ctx->watchdog_s = env_int("FMPNP_WATCHDOG_S", 20);
    if (ctx->watchdog_s < 0) ctx->watchdog_s = 0;
    ctx->ctas_per_sm = env_int("FMPNP_CTAS_PER_SM", 0);
    ctx->reuse = env_int("FMPNP_REUSE", 1);
    ctx->light = env_int("FMPNP_LIGHT", -1);
    ctx->l2_policy = env_int("FMPNP_L2_POLICY", 1) != 0 ? 1 : 0;
    ctx->light_miss_den = env_int("FMPNP_LIGHT_MISS_DEN", 64);
    if (ctx->light_miss_den < 1) ctx->light_miss_den = 1;
    {  // fmpnp_compute_cost_batch_noalloc: resolve its scratch array now, so that the call itself only launches
        void* scratch = nullptr;
        if (static_scratch_setup(device, kNoallocRingBytes, &scratch) != cudaSuccess) cudaGetLastError();  // reported by the call
    }
    *out = ctx;
    return FMPNP_SUCCESS;
}

int fmpnp_destroy(fmpnp_ctx* ctx) {
    if (!ctx) return FMPNP_SUCCESS;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    if (ctx->d_queries) cudaFree(ctx->d_queries);
    for (int i = 0; i < 2; ++i)
        if (ctx->h_queries[i]) cudaFreeHost(ctx->h_queries[i]);
    if (ctx->d_rec) cudaFree(ctx->d_rec);
    if (ctx->d_supp) cudaFree(ctx->d_supp);
    if (ctx->d_inl) cudaFree(ctx->d_inl);
    if (ctx->d_counter) cudaFree(ctx->d_counter);
    if (ctx->d_states) cudaFree(ctx->d_states);
    if (ctx->d_part) cudaFree(ctx->d_part);
    if (ctx->d_queue) cudaFree(ctx->d_queue);
    if (ctx->d_sched) cudaFree(ctx->d_sched);
    for (int i = 0; i < 2; ++i)
        if (ctx->staged[i]) cudaEventDestroy(ctx->staged[i]);
    if (ctx->last_done) cudaEventDestroy(ctx->last_done);
    delete ctx;
    return FMPNP_SUCCESS;
}

const char* fmpnp_last_error(const fmpnp_ctx* ctx) { return ctx ? ctx->err : "null context"; }

int64_t fmpnp_launch_count(const fmpnp_ctx* ctx) { return ctx ? ctx->launches : 0; }

int fmpnp_last_geometry(const fmpnp_ctx* ctx, int32_t* grid, int32_t* cluster_size, int32_t* ring_bytes,
                        int32_t* tasks_per_evaluation) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (tasks_per_evaluation) *tasks_per_evaluation = ctx->last.ctas;
    if (grid) *grid = ctx->last.grid;
    if (cluster_size) *cluster_size = ctx->last.cluster;
    if (ring_bytes) *ring_bytes = ctx->last.ring_bytes;
    return FMPNP_SUCCESS;
}

int fmpnp_refine_batch(fmpnp_ctx* ctx, const fmpnp_problem* probs, int32_t B, const fmpnp_options* o,
                       fmpnp_result* results, const fmpnp_history* hist, void* stream_) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (B == 0) return FMPNP_SUCCESS;
    if (!probs || !o || !results || B < 0) return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "null argument or B < 0");
    if (o->n_iters < 0) return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "n_iters < 0");
    if (o->loss_id < FMPNP_LOSS_SQUARED || o->loss_id > FMPNP_LOSS_BARRON)
        return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "unknown loss id %d", o->loss_id);
    if (o->n_levels < 0 || o->n_levels > FMPNP_MAX_LEVELS)
        return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "n_levels must be in [0, %d]", FMPNP_MAX_LEVELS);
    if (o->solver != FMPNP_SOLVER_CHOLESKY && o->solver != FMPNP_SOLVER_LU)
        return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "unknown solver %d", o->solver);
    for (int32_t b = 0; b < B; ++b) {
        int rc = validate_problem(ctx, probs[b], b, true);
        if (rc) return rc;
        for (int l = 0; l < o->n_levels; ++l)
            if (o->levels[l][0] < 0 || o->levels[l][1] > probs[b].C || o->levels[l][0] >= o->levels[l][1])
                return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "level %d = [%d, %d) outside the %d channels of problem %d", l,
                            o->levels[l][0], o->levels[l][1], probs[b].C, b);
    }
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    FMPNP_CUDA(ctx, cudaSetDevice(ctx->device));
    if (int rc0 = order_after_previous(ctx, stream)) return rc0;
    Geometry g;
    int rc = geometry(ctx, B, max_points(probs, B), o->cluster_size, &g, ctx->reuse != 0 && o->gather_every_row == 0, !std::isnan(o->ratio_threshold));
    if (rc) return rc;
    int32_t n_max;
    rc = stage(ctx, probs, B, g.units > 0 ? B : g.grid, &n_max, stream);
    if (rc) return rc;
    const int levels_eff = o->n_levels > 0 ? o->n_levels : 1;
    if (hist) {
        if (hist->cap < levels_eff * (o->n_iters + 1) || hist->n_stride < n_max || !hist->len || !hist->Rs || !hist->ts ||
            !hist->costs || !hist->points2d || !hist->mask || !hist->kept)
            return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "history buffers too small or null (cap %d, n_stride %d)", hist->cap,
                        hist->n_stride);
    }
    LaunchParams lp;
    fill_common(lp, ctx, B, n_max, g);
    rc = prepare_tasks(ctx, B, g, lp, stream);
    if (rc) return rc;
    lp.mode = kModeRefine;
    lp.n_iters = o->n_iters;
    lp.loss_id = o->loss_id;
    lp.alpha = o->barron_alpha;
    lp.lambda0 = o->lambda0;
    lp.use_ratio = std::isnan(o->ratio_threshold) ? 0 : 1;
    lp.ratio_thr = lp.use_ratio ? o->ratio_threshold : 0.0;
    lp.n_levels = o->n_levels;
    memcpy(lp.levels, o->levels, sizeof(lp.levels));
    lp.solver = o->solver;
    lp.results = results;
    lp.reuse = (ctx->reuse != 0 && o->gather_every_row == 0) ? 1 : 0;
    lp.light = lp.reuse && g.cluster == 1 && light_route(ctx, B, lp.use_ratio != 0) ? 1 : 0;
    lp.inlier_on = (o->inlier_threshold == o->inlier_threshold && o->inlier_threshold != 0.0) ? 1 : 0;  // 0 / NaN = off
    lp.inlier_thr = o->inlier_threshold;
    // fmpnp_options.keep_records: the same arrays as the previous refinement launch of this context (checked: pointers,
    // sizes, batch, record layout) and nothing in between -> that launch's generations again, so its records match
    const QueryDev* staged_now = ctx->h_queries[ctx->stage_slot ^ 1];  // what stage() has just filled
    bool keep = o->keep_records != 0 && lp.reuse != 0 && o->n_levels == 0 && g.units > 0 && ctx->prev_valid &&
                ctx->prev_B == B && ctx->prev_n_max == n_max && (int32_t)ctx->prev_queries.size() == B;
    for (int32_t b = 0; keep && b < B; ++b) {
        const QueryDev &x = staged_now[b], &y = ctx->prev_queries[b];
        keep = x.F == y.F && x.Gx == y.Gx && x.Gy == y.Gy && x.ref == y.ref && x.pts == y.pts && x.N == y.N && x.C == y.C &&
               x.Hf == y.Hf && x.Wf == y.Wf && x.im_w == y.im_w && x.im_h == y.im_h;
    }
    if (keep) {
        lp.reuse_base = ctx->prev_base;
    } else {  // one generation per (query, level) of this launch; generations are never handed out twice
        const uint64_t gens = (uint64_t)B * (uint64_t)levels_eff;
        if (ctx->reuse_next + gens >= (1ull << 32)) {
            FMPNP_CUDA(ctx, cudaMemsetAsync(ctx->d_rec, 0, ctx->cap_points * kRecDoubles * sizeof(double), stream));
            ctx->reuse_next = 1;
        }
        if (gens >= (1ull << 32) - 1) lp.reuse = 0;
        lp.reuse_base = (uint32_t)ctx->reuse_next;
        ctx->reuse_next += gens;
    }
    ctx->prev_valid = lp.reuse != 0 && o->n_levels == 0 && g.units > 0;
    if (ctx->prev_valid) {
        ctx->prev_B = B;
        ctx->prev_n_max = n_max;
        ctx->prev_base = lp.reuse_base;
        try {
            ctx->prev_queries.assign(staged_now, staged_now + B);
        } catch (...) {
            ctx->prev_valid = false;
        }
    }
    if (hist) {
        lp.hist.enabled = 1;
        lp.hist.cap = hist->cap;
        lp.hist.n_stride = hist->n_stride;
        lp.hist.len = hist->len;
        lp.hist.Rs = hist->Rs;
        lp.hist.ts = hist->ts;
        lp.hist.costs = hist->costs;
        lp.hist.points2d = hist->points2d;
        lp.hist.mask = hist->mask;
        lp.hist.kept = hist->kept;
    }
    FMPNP_CUDA(ctx, cudaMemsetAsync(ctx->d_counter, 0, sizeof(unsigned int), stream));
    FMPNP_CUDA(ctx, launch_refine(lp, g.grid, g.ring_bytes, stream));
    if (int rc1 = mark_launched(ctx, stream)) return rc1;
    ctx->launches += 1;
    ctx->last = {g.grid, g.cluster, g.ring_bytes, g.units};
    return FMPNP_SUCCESS;
}

int fmpnp_compute_cost_batch(fmpnp_ctx* ctx, const fmpnp_problem* probs, int32_t B, const fmpnp_options* o, double* costs,
                             int32_t* n_supported, void* stream_) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (B == 0) return FMPNP_SUCCESS;
    if (!probs || !o || !costs || B < 0) return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "null argument or B < 0");
    for (int32_t b = 0; b < B; ++b) {
        int rc = validate_problem(ctx, probs[b], b, false);
        if (rc) return rc;
    }
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    FMPNP_CUDA(ctx, cudaSetDevice(ctx->device));
    if (int rc0 = order_after_previous(ctx, stream)) return rc0;
    Geometry g;
    int rc = geometry(ctx, B, max_points(probs, B), 0, &g);
    if (rc) return rc;
    int32_t n_max;
    rc = stage(ctx, probs, B, g.units > 0 ? B : g.grid, &n_max, stream);
    if (rc) return rc;
    LaunchParams lp;
    fill_common(lp, ctx, B, n_max, g);
    rc = prepare_tasks(ctx, B, g, lp, stream);
    if (rc) return rc;
    ctx->prev_valid = false;  // this launch overwrites records (fmpnp_options.keep_records)
    lp.mode = kModeComputeCost;
    lp.use_ratio = std::isnan(o->ratio_threshold) ? 0 : 1;
    lp.ratio_thr = lp.use_ratio ? o->ratio_threshold : 0.0;
    lp.cost_out = costs;
    lp.nsup_out = n_supported;
    FMPNP_CUDA(ctx, cudaMemsetAsync(ctx->d_counter, 0, sizeof(unsigned int), stream));
    FMPNP_CUDA(ctx, launch_refine(lp, g.grid, g.ring_bytes, stream));
    if (int rc1 = mark_launched(ctx, stream)) return rc1;
    ctx->launches += 1;
    ctx->last = {g.grid, g.cluster, g.ring_bytes, g.units};
    return FMPNP_SUCCESS;
}

// No counterpart in the reference: compute_cost for callers that cannot allocate (see include/fmpnp.h).  Calls no
// cudaMalloc / cudaFree / cudaMallocHost and touches none of the context's workspace: the descriptors are a kernel
// parameter, the scratch is the library's file-scope device array (fmpnp_kernels.cu).
int fmpnp_compute_cost_batch_noalloc(fmpnp_ctx* ctx, const fmpnp_problem* probs, int32_t B, const fmpnp_options* o,
                                     double* costs, int32_t* n_supported, void* stream_) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (B == 0) return FMPNP_SUCCESS;
    if (!probs || !o || !costs || B < 0) return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "null argument or B < 0");
    if (B > FMPNP_NOALLOC_MAX_QUERIES)
        return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT,
                    "fmpnp_compute_cost_batch_noalloc: %d queries exceed the static scratch (at most %d per call; use "
                    "fmpnp_compute_cost_batch)", B, FMPNP_NOALLOC_MAX_QUERIES);
    StaticQueries pack;
    memset(&pack, 0, sizeof(pack));
    for (int32_t b = 0; b < B; ++b) {
        const fmpnp_problem& p = probs[b];
        int rc = validate_problem(ctx, p, b, false);
        if (rc) return rc;
        if (p.N > FMPNP_NOALLOC_MAX_POINTS)
            return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT,
                        "fmpnp_compute_cost_batch_noalloc: problem %d has %d points, the static scratch holds at most %d "
                        "per query (use fmpnp_compute_cost_batch)", b, p.N, FMPNP_NOALLOC_MAX_POINTS);
        QueryDev& q = pack.q[b];
        q.F = p.F;
        q.Gx = p.Gx;
        q.Gy = p.Gy;
        q.ref = p.ref;
        q.pts = p.pts3D;
        q.active = p.active;
        memcpy(q.K, p.K, sizeof(q.K));
        memcpy(q.R0, p.R0, sizeof(q.R0));
        memcpy(q.t0, p.t0, sizeof(q.t0));
        q.N = p.N;
        q.C = p.C;
        q.Hf = p.Hf;
        q.Wf = p.Wf;
        q.im_w = p.im_width;
        q.im_h = p.im_height;
    }
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    FMPNP_CUDA(ctx, cudaSetDevice(ctx->device));
    void* scratch = nullptr;
    FMPNP_CUDA(ctx, static_scratch_setup(ctx->device, kNoallocRingBytes, &scratch));
    LaunchParams lp;
    memset(&lp, 0, sizeof(lp));
    lp.B = B;
    lp.mode = kModeComputeCost;
    lp.ring_per_warp = kNoallocRingBytes / kWarps;
    lp.min_depth = 2;
    lp.cluster_size = 1;  // one CTA per query, no task queue: nothing but records, flags and the counter to keep
    lp.scheduler = 0;
    lp.use_ratio = std::isnan(o->ratio_threshold) ? 0 : 1;
    lp.ratio_thr = lp.use_ratio ? o->ratio_threshold : 0.0;
    lp.cost_out = costs;
    lp.nsup_out = n_supported;
    FMPNP_CUDA(ctx, launch_compute_cost_static(lp, pack, scratch, B, kNoallocRingBytes, stream));
    ctx->launches += 1;
    return FMPNP_SUCCESS;
}

int fmpnp_find_inliers_batch(fmpnp_ctx* ctx, const fmpnp_problem* probs, int32_t B, int32_t loss_id, double barron_alpha,
                             double threshold, uint8_t* inliers, int32_t n_stride, void* stream_) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (B == 0) return FMPNP_SUCCESS;
    if (!probs || !inliers || B < 0) return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "null argument or B < 0");
    if (loss_id < FMPNP_LOSS_SQUARED || loss_id > FMPNP_LOSS_BARRON)
        return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "unknown loss id %d", loss_id);
    for (int32_t b = 0; b < B; ++b) {
        int rc = validate_problem(ctx, probs[b], b, false);
        if (rc) return rc;
        if (probs[b].N > n_stride) return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "n_stride %d < N of problem %d", n_stride, b);
    }
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    FMPNP_CUDA(ctx, cudaSetDevice(ctx->device));
    if (int rc0 = order_after_previous(ctx, stream)) return rc0;
    Geometry g;
    int rc = geometry(ctx, B, max_points(probs, B), 0, &g);
    if (rc) return rc;
    int32_t n_max;
    rc = stage(ctx, probs, B, g.units > 0 ? B : g.grid, &n_max, stream);
    if (rc) return rc;
    LaunchParams lp;
    fill_common(lp, ctx, B, n_max, g);
    rc = prepare_tasks(ctx, B, g, lp, stream);
    if (rc) return rc;
    ctx->prev_valid = false;  // this launch overwrites records (fmpnp_options.keep_records)
    lp.mode = kModeFindInliers;
    lp.loss_id = loss_id;
    lp.alpha = barron_alpha;
    lp.use_ratio = 1;
    lp.ratio_thr = threshold;
    lp.inliers_out = inliers;
    lp.inliers_stride = n_stride;
    FMPNP_CUDA(ctx, cudaMemsetAsync(ctx->d_counter, 0, sizeof(unsigned int), stream));
    FMPNP_CUDA(ctx, launch_refine(lp, g.grid, g.ring_bytes, stream));
    if (int rc1 = mark_launched(ctx, stream)) return rc1;
    ctx->launches += 1;
    ctx->last = {g.grid, g.cluster, g.ring_bytes, g.units};
    return FMPNP_SUCCESS;
}

int fmpnp_pack_chw_to_hwc(fmpnp_ctx* ctx, const void* src, int32_t src_is_f64, float* dst, int32_t C, int32_t Hf,
                          int32_t Wf, void* stream_) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (!src || !dst || C < 1 || Hf < 1 || Wf < 1) return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "bad pack arguments");
    FMPNP_CUDA(ctx, cudaSetDevice(ctx->device));
    FMPNP_CUDA(ctx, launch_pack_chw_to_hwc(src, src_is_f64, dst, C, Hf, Wf, static_cast<cudaStream_t>(stream_)));
    ctx->launches += 1;
    return FMPNP_SUCCESS;
}

int fmpnp_sobel_pack(fmpnp_ctx* ctx, const void* src, int32_t src_is_f64, float* F, float* Gx, float* Gy, int32_t C,
                     int32_t Hf, int32_t Wf, int32_t border, int32_t normalized, void* stream_) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (!src || !F || !Gx || !Gy || C < 1 || Hf < 1 || Wf < 1 ||
        (border != FMPNP_BORDER_REPLICATE && border != FMPNP_BORDER_ZERO))
        return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "bad sobel_pack arguments");
    FMPNP_CUDA(ctx, cudaSetDevice(ctx->device));
    FMPNP_CUDA(ctx, launch_sobel_pack(src, src_is_f64, F, Gx, Gy, C, Hf, Wf, border == FMPNP_BORDER_ZERO, normalized,
                                      static_cast<cudaStream_t>(stream_)));
    ctx->launches += 1;
    return FMPNP_SUCCESS;
}

int fmpnp_gather_descriptors(fmpnp_ctx* ctx, const void* src, int32_t src_is_f64, const int32_t* rows, const int32_t* cols,
                             float* dst, int32_t C, int32_t H, int32_t W, int32_t N, void* stream_) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (N == 0) return FMPNP_SUCCESS;
    if (!src || !rows || !cols || !dst || C < 1 || H < 1 || W < 1 || N < 0)
        return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "bad gather_descriptors arguments");
    FMPNP_CUDA(ctx, cudaSetDevice(ctx->device));
    FMPNP_CUDA(ctx, launch_gather_chw_rows(src, src_is_f64, rows, cols, dst, C, H, W, N, static_cast<cudaStream_t>(stream_)));
    ctx->launches += 1;
    return FMPNP_SUCCESS;
}

int fmpnp_resize_bilinear_chw(fmpnp_ctx* ctx, const void* src, int32_t src_is_f64, double* dst, int32_t C, int32_t H,
                              int32_t W, int32_t Ho, int32_t Wo, void* stream_) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (!src || !dst || C < 1 || H < 1 || W < 1 || Ho < 1 || Wo < 1)
        return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "bad resize arguments");
    FMPNP_CUDA(ctx, cudaSetDevice(ctx->device));
    FMPNP_CUDA(ctx, launch_bilinear_resize_chw(src, src_is_f64, dst, C, H, W, Ho, Wo, static_cast<cudaStream_t>(stream_)));
    ctx->launches += 1;
    return FMPNP_SUCCESS;
}

int fmpnp_blur_chw(fmpnp_ctx* ctx, const void* src, int32_t src_is_f64, const double* kernel, double* dst, int32_t C,
                   int32_t H, int32_t W, int32_t k, int32_t pad, void* stream_) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (!src || !kernel || !dst || C < 1 || H < 1 || W < 1 || k < 1 || pad < 0 || H + 2 * pad - k + 1 < 1 ||
        W + 2 * pad - k + 1 < 1)
        return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "bad blur arguments");
    FMPNP_CUDA(ctx, cudaSetDevice(ctx->device));
    FMPNP_CUDA(ctx, launch_blur_chw(src, src_is_f64, kernel, dst, C, H, W, k, pad, static_cast<cudaStream_t>(stream_)));
    ctx->launches += 1;
    return FMPNP_SUCCESS;
}

// ---- result records in another GPU's memory (multi-GPU gather without a collective) ----------------------
int fmpnp_shared_buffer_create(fmpnp_ctx* ctx, size_t bytes, void** dev_ptr, unsigned char* handle) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (!dev_ptr || !handle || bytes == 0) return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "bad shared_buffer_create arguments");
    static_assert(sizeof(cudaIpcMemHandle_t) == FMPNP_IPC_HANDLE_BYTES, "IPC handle size");
    FMPNP_CUDA(ctx, cudaSetDevice(ctx->device));
    void* p = nullptr;
    FMPNP_CUDA(ctx, cudaMalloc(&p, bytes));
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaMemset(p, 0, bytes);
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        return fail(ctx, FMPNP_ERR_CUDA, "cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    }
    memcpy(handle, &h, sizeof(h));
    *dev_ptr = p;
    return FMPNP_SUCCESS;
}

int fmpnp_shared_buffer_open(fmpnp_ctx* ctx, const unsigned char* handle, void** dev_ptr) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (!dev_ptr || !handle) return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "bad shared_buffer_open arguments");
    FMPNP_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    FMPNP_CUDA(ctx, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *dev_ptr = p;
    return FMPNP_SUCCESS;
}

int fmpnp_shared_buffer_close(fmpnp_ctx* ctx, void* dev_ptr) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (!dev_ptr) return FMPNP_SUCCESS;
    FMPNP_CUDA(ctx, cudaSetDevice(ctx->device));
    FMPNP_CUDA(ctx, cudaIpcCloseMemHandle(dev_ptr));
    return FMPNP_SUCCESS;
}

int fmpnp_shared_buffer_destroy(fmpnp_ctx* ctx, void* dev_ptr) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (!dev_ptr) return FMPNP_SUCCESS;
    FMPNP_CUDA(ctx, cudaSetDevice(ctx->device));
    FMPNP_CUDA(ctx, cudaFree(dev_ptr));
    return FMPNP_SUCCESS;
}

int fmpnp_copy_to_host(fmpnp_ctx* ctx, void* dst_host, const void* src_dev, size_t bytes, void* stream_) {
    if (!ctx) return FMPNP_ERR_INVALID_ARGUMENT;
    if (bytes == 0) return FMPNP_SUCCESS;
    if (!dst_host || !src_dev) return fail(ctx, FMPNP_ERR_INVALID_ARGUMENT, "bad copy_to_host arguments");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    FMPNP_CUDA(ctx, cudaSetDevice(ctx->device));
    FMPNP_CUDA(ctx, cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, stream));
    FMPNP_CUDA(ctx, cudaStreamSynchronize(stream));
    return FMPNP_SUCCESS;
}

}  // extern "C"
