// libgnss_b200.so: host side of the C ABI declared in include/gnss_b200.h.
// One gb200_service owns the CUDA stream, the per-stream device sample rings, the FFT plans and the
// scratch pools; acquisition and tracking handles borrow them.  No CPU compute path exists here: if no
// sm_100 device is present, gb200_service_create fails and nothing else can be called.
#include "gb200_internal.cuh"
#include <algorithm>
#include <sched.h>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <string>
#include <vector>

#define GB_CUDA(s, call)                                                                                  \
    do                                                                                                    \
        {                                                                                                 \
            cudaError_t e__ = (call);                                                                     \
            if (e__ != cudaSuccess)                                                                       \
                {                                                                                         \
                    (s)->set_error(std::string(#call) + ": " + cudaGetErrorString(e__));                  \
                    return GB200_E_CUDA;                                                                  \
                }                                                                                         \
        }                                                                                                 \
    while (0)

namespace
{
std::string g_global_error = "no error";
}

struct GbRing
{
    void* d_buf = nullptr;
    int item_type = -1;
    uint64_t cap = 0, slack = 0;
    uint64_t end = 0;    // one past the newest item index held
    uint64_t begin = 0;  // oldest item index still held
    bool any = false;
    bool conj = false;   // inverted_spectrum: items are conjugated as they are pushed
    cudaEvent_t ev_pushed = nullptr;  // recorded on the copy stream behind the newest push
    bool pushed_unseen = false;       // the compute stream has not been ordered behind ev_pushed yet
    // the last GB_RING_MARKS pushes, each with its own event: a tracking request waits only for the push that delivered its
    // last sample, so one call over several pushed blocks starts on the early blocks while the later ones still land
    struct Mark
    {
        uint64_t end = 0;   // one past the newest item of that push (0: slot unused)
        uint64_t seq = 0;   // push order
        cudaEvent_t ev = nullptr;
    };
    static const int kMarks = 16;
    Mark marks[kMarks];
    uint64_t push_seq = 0;
    // acquisition resampler hanging off this (source) stream: gb200_stream_set_decimator
    struct Decimator
    {
        bool on = false;
        int dst = -1;
        unsigned D = 0, T = 0;
        float* d_taps = nullptr;
        float2* d_out = nullptr;     // one push worth of outputs, then copied into the destination ring in stream order
        size_t out_cap = 0;
        uint64_t next_k = 0;         // next output index to produce
        uint64_t valid_begin = 0;    // first source item of the current continuous run (older ones count as zero history)
    } decim;
    // event of the oldest recorded push that already covers item index `need_end - 1` (nullptr: none recorded)
    const Mark* mark_for(uint64_t need_end) const
    {
        const Mark* best = nullptr;
        for (int i = 0; i < kMarks; i++)
            if (marks[i].end >= need_end && marks[i].end != 0 && (!best || marks[i].seq < best->seq)) best = &marks[i];
        return best;
    }
};

struct GbPlanEntry
{
    GbFftPlan plan;
    float2* d_tw1 = nullptr;
    float2* d_tw2 = nullptr;
    float2* d_twN = nullptr;
    float2* d_twNhi = nullptr;
};

// optional per-kernel-class CUDA-event timing (bench roofline); off by default
struct GbKernelTiming
{
    bool enabled = false;
    struct Pending { int cls; cudaEvent_t a, b; };
    std::vector<Pending> pending;
    std::vector<cudaEvent_t> pool;
    double ms[GB200_KCLASS_COUNT] = {0};
    uint64_t count[GB200_KCLASS_COUNT] = {0};
};

// ---- batched acquisition: launch -> collect hand-over (one per service) ----
struct AcqGroup
{
    std::vector<int> members;  // indices into the caller's list
    std::vector<float> steps;
    bool grid = false;
    int mode_key = 0;          // grid-keeping handles share a group only when they run the same column-kernel mode
    size_t off_segs = 0, off_segoff = 0, off_code = 0, off_grid = 0, off_tmp = 0;  // byte offsets in the upload arena
    size_t off_power = 0, off_peaks = 0, off_second = 0;                          // byte offsets in the result arena
};
struct AcqBatchCtx
{
    std::vector<AcqGroup> groups;
    bool active = false;
    // results land in a pinned buffer of the batch's own (not the service's shared staging area, which
    // gb200_acq_set_local_code reuses between launch and collect)
    char* h_res = nullptr;
    size_t h_res_bytes = 0;
    size_t res_bytes = 0;
    // signature of the launched batch: when the next launch has the same one the group tables on the device are reused
    std::vector<uint64_t> sig;
    size_t up_bytes = 0;
};

// Tuning switches of the acquisition inverse side (gb200_service_set_tuning); 0 = the default of each
struct GbTuning
{
    int lanes = 2;        // streams the inverse chunks alternate on
    int l2_mb = 48;       // row-buffer budget across the lanes
    int chunk_t = 0;      // transforms per chunk (0: from l2_mb)
    int fused = 0;        // one persistent launch for the inverse side
    int fused_lag = 80, fused_ring = 176;
    int fwd_groups = 1;   // bin groups of the forward side; > 1: bin-major transform order, inverse chunks start after the first group
                          // (measured: no gain at C2 / C5, a loss on small batches -- profiles/r02_acq_fwd_sweep.txt)
    int split_n1 = 0;     // force N1 of the four-step split
    int variant = 0;      // alternative kernel instantiation of a split
    int generic = 0;      // run-time parametrised kernels only
    int fwd_generic = 0;  // run-time parametrised kernels on the forward side only
    int host_threads = 8; // request staging threads of long tracking batches
    int trk_chunk_max = 32768;
    int trace = 0;
};

struct gb200_service
{
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;  // second lane for the inverse chunks (fills the tail of the other lane's kernels)
    cudaStream_t stream34[2] = {nullptr, nullptr};  // optional third / fourth lane (GB200_ACQ_LANES)
    cudaEvent_t ev_join34[2] = {nullptr, nullptr};
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaEvent_t ev_fwd[8] = {};  // one per forward bin group (recorded on the forward stream, stream34[0])
    cudaEvent_t ev_power_fork = nullptr;  // the input-power reduction runs on stream2 beside the forward transforms
    // Ingest runs on its own stream so that the H2D copy of block k+1 overlaps the kernels working on block k.  Ordering:
    // a consumer launch waits for the newest push of every ring (ev_pushed); a push waits for the newest consumer launch
    // (ev_consumed) before it may overwrite ring memory.
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_consumed = nullptr;
    bool any_consumed = false;
    std::vector<GbRing> rings;
    uint64_t ring_capacity = 0, ring_slack = 0;
    std::map<int, GbPlanEntry> plans;
    uint64_t launches = 0;
    // coalescing of concurrent single-request callers ("the service coalesces concurrent callers into one launch")
    struct TrkPending
    {
        gb200_trk_req req;
        float* out;
        int rc;
        bool done;
    };
    uint32_t batch_window_us = 0;
    std::mutex qmu;
    std::condition_variable qcv;
    std::vector<TrkPending*> queue;
    bool collecting = false;
    bool running = false;   // a coalesced batch is executing: the next collector keeps collecting meanwhile (it could not launch anyway)
    int live_trk = 0;       // float-correlator handles alive
    int prev_batch_n = 1;   // callers served by the previous coalesced launch: the collector waits for that many (adaptive)
    uint64_t coalesced_batches = 0, coalesced_requests = 0;
    // the same for concurrent gb200_acq_run callers (the channels' acquisitions over one sample block become
    // ONE gb200_acq_run_batch, so the Doppler-wiped forward spectra are computed once for all of them)
    struct AcqPending
    {
        gb200_acq* a;
        int stream_id;
        uint64_t first;
        gb200_acq_result* out;
        int rc;
        bool done;
    };
    std::mutex aqmu;
    std::condition_variable aqcv;
    std::vector<AcqPending*> acq_queue;
    bool acq_collecting = false;
    uint64_t acq_coalesced_batches = 0, acq_coalesced_requests = 0;
    gb200_trk_batch* cb = nullptr;  // arena of gb200_trk_correlate_batch (grow-only)
    int cb_n = 0, cb_chunks = 0;
    std::string error = "no error";
    std::recursive_mutex mu;
    // scratch pools (grown on demand, reused)
    void* d_scratch[20] = {};
    size_t scratch_bytes[20] = {};
    void* h16 = nullptr;  // pinned staging of the int16 correlator (work items in, results out)
    size_t h16_bytes = 0;
    void* h_pinned = nullptr;
    size_t h_pinned_bytes = 0;
    int sm_count = 148;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    GbKernelTiming kt;
    void* d_flush = nullptr;
    size_t flush_bytes = 0;
    AcqBatchCtx acq_ctx;
    GbTuning tune;

    void set_error(const std::string& e)
    {
        static std::mutex emu;
        std::lock_guard<std::mutex> lk(emu);
        error = e;
        g_global_error = e;
    }
    int scratch(int slot, size_t bytes, void** out)
    {
        if (bytes > scratch_bytes[slot])
            {
                if (d_scratch[slot]) cudaFree(d_scratch[slot]);
                d_scratch[slot] = nullptr;
                scratch_bytes[slot] = 0;
                size_t want = bytes + bytes / 4;
                cudaError_t e = cudaMalloc(&d_scratch[slot], want);
                if (e != cudaSuccess)
                    {
                        set_error(std::string("cudaMalloc scratch: ") + cudaGetErrorString(e));
                        return GB200_E_CUDA;
                    }
                scratch_bytes[slot] = want;
            }
        *out = d_scratch[slot];
        return GB200_OK;
    }
    int pinned(size_t bytes, void** out)
    {
        if (bytes > h_pinned_bytes)
            {
                if (h_pinned) cudaFreeHost(h_pinned);
                h_pinned = nullptr;
                h_pinned_bytes = 0;
                cudaError_t e = cudaMallocHost(&h_pinned, bytes * 2);
                if (e != cudaSuccess)
                    {
                        set_error(std::string("cudaMallocHost: ") + cudaGetErrorString(e));
                        return GB200_E_CUDA;
                    }
                h_pinned_bytes = bytes * 2;
            }
        *out = h_pinned;
        return GB200_OK;
    }
};

extern "C" const char* gb200_version(void) { return "gnss_b200 0.1 (sm_100a)"; }

extern "C" const char* gb200_last_error(const gb200_service* s) { return s ? s->error.c_str() : g_global_error.c_str(); }

extern "C" int gb200_service_create(const gb200_service_cfg* cfg, gb200_service** out)
{
    if (!out) return GB200_E_INVALID;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        {
            g_global_error = "no CUDA device: libgnss_b200 has no CPU fallback";
            return GB200_E_NO_DEVICE;
        }
    int dev = (cfg && cfg->device >= 0) ? cfg->device : 0;
    if (cfg && cfg->device < 0) cudaGetDevice(&dev);
    if (dev >= count)
        {
            g_global_error = "device ordinal out of range";
            return GB200_E_INVALID;
        }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess || prop.major != 10)
        {
            g_global_error = "device is not sm_100 (Blackwell B200): the kernels are built for sm_100a only";
            return GB200_E_NO_DEVICE;
        }
    if (cudaSetDevice(dev) != cudaSuccess)
        {
            g_global_error = "cudaSetDevice failed";
            return GB200_E_CUDA;
        }
    gb200_service* s = new gb200_service();
    s->device = dev;
    s->sm_count = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&s->stream2, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&s->stream34[0], cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&s->stream34[1], cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&s->ev_join34[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&s->ev_join34[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&s->ev_consumed, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&s->ev_power_fork, cudaEventDisableTiming) != cudaSuccess ||
        ![&] { for (auto& e : s->ev_fwd) if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return false; return true; }() ||
        cudaEventCreateWithFlags(&s->ev_join, cudaEventDisableTiming) != cudaSuccess)
        {
            g_global_error = "cudaStreamCreate failed";
            delete s;
            return GB200_E_CUDA;
        }
    s->rings.resize((cfg && cfg->max_streams > 0) ? cfg->max_streams : 4);
    s->ring_capacity = (cfg && cfg->ring_capacity) ? cfg->ring_capacity : (1ull << 24);
    s->ring_slack = (cfg && cfg->ring_slack) ? cfg->ring_slack : (1ull << 21);
    s->batch_window_us = cfg ? cfg->batch_window_us : 0;
    if (s->ring_slack > s->ring_capacity) s->ring_slack = s->ring_capacity;
    *out = s;
    return GB200_OK;
}

extern "C" int gb200_service_destroy(gb200_service* s)
{
    if (!s) return GB200_E_INVALID;
    cudaSetDevice(s->device);
    if (s->copy_stream) cudaStreamSynchronize(s->copy_stream);
    cudaStreamSynchronize(s->stream);
    if (s->cb) gb200_trk_batch_destroy(s->cb);
    for (auto& r : s->rings)
        {
            if (r.d_buf) cudaFree(r.d_buf);
            if (r.decim.d_taps) cudaFree(r.decim.d_taps);
            if (r.decim.d_out) cudaFree(r.decim.d_out);
        }
    for (auto& kv : s->plans)
        {
            if (kv.second.d_tw1) cudaFree(kv.second.d_tw1);
            if (kv.second.d_tw2) cudaFree(kv.second.d_tw2);
            if (kv.second.d_twN) cudaFree(kv.second.d_twN);
            if (kv.second.d_twNhi) cudaFree(kv.second.d_twNhi);
        }
    for (int i = 0; i < 20; i++)
        if (s->d_scratch[i]) cudaFree(s->d_scratch[i]);
    if (s->h16) cudaFreeHost(s->h16);
    if (s->h_pinned) cudaFreeHost(s->h_pinned);
    if (s->acq_ctx.h_res) cudaFreeHost(s->acq_ctx.h_res);
    if (s->d_flush) cudaFree(s->d_flush);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->ev_fork) cudaEventDestroy(s->ev_fork);
    if (s->ev_power_fork) cudaEventDestroy(s->ev_power_fork);
    for (auto& e : s->ev_fwd) if (e) cudaEventDestroy(e);
    if (s->ev_join) cudaEventDestroy(s->ev_join);
    if (s->stream2) cudaStreamDestroy(s->stream2);
    for (int k = 0; k < 2; k++)
        {
            if (s->ev_join34[k]) cudaEventDestroy(s->ev_join34[k]);
            if (s->stream34[k]) cudaStreamDestroy(s->stream34[k]);
        }
    if (s->ev_consumed) cudaEventDestroy(s->ev_consumed);
    if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
    for (auto& r : s->rings)
        {
            if (r.ev_pushed) cudaEventDestroy(r.ev_pushed);
            for (auto& m : r.marks)
                if (m.ev) cudaEventDestroy(m.ev);
        }
    cudaStreamDestroy(s->stream);
    delete s;
    return GB200_OK;
}

extern "C" void* gb200_service_cuda_stream(gb200_service* s) { return s ? (void*)s->stream : nullptr; }

extern "C" int gb200_service_synchronize(gb200_service* s)
{
    if (!s) return GB200_E_INVALID;
    GB_CUDA(s, cudaStreamSynchronize(s->copy_stream));
    GB_CUDA(s, cudaStreamSynchronize(s->stream));
    return GB200_OK;
}

// order the compute stream behind every push issued so far (call before launching anything that reads a ring)
static int rings_acquire(gb200_service* s)
{
    for (auto& r : s->rings)
        if (r.pushed_unseen)
            {
                GB_CUDA(s, cudaStreamWaitEvent(s->stream, r.ev_pushed, 0));
                r.pushed_unseen = false;
            }
    return GB200_OK;
}

// mark the kernels launched so far as readers of ring memory (call after launching them)
static int rings_release(gb200_service* s)
{
    GB_CUDA(s, cudaEventRecord(s->ev_consumed, s->stream));
    s->any_consumed = true;
    return GB200_OK;
}

extern "C" uint64_t gb200_service_launch_count(const gb200_service* s) { return s ? s->launches : 0; }

extern "C" int gb200_service_timer_start(gb200_service* s)
{
    if (!s) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    if (!s->ev0)
        {
            GB_CUDA(s, cudaEventCreate(&s->ev0));
            GB_CUDA(s, cudaEventCreate(&s->ev1));
        }
    GB_CUDA(s, cudaEventRecord(s->ev0, s->stream));
    GB_CUDA(s, cudaStreamWaitEvent(s->copy_stream, s->ev0, 0));  // pushes issued after the start mark begin after it
    return GB200_OK;
}

extern "C" int gb200_service_timer_stop(gb200_service* s, float* elapsed_ms)
{
    if (!s || !elapsed_ms || !s->ev0) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    { int rc = rings_acquire(s); if (rc) return rc; }  // a push with no consumer behind it still ends inside the timed region
    GB_CUDA(s, cudaEventRecord(s->ev1, s->stream));
    GB_CUDA(s, cudaEventSynchronize(s->ev1));
    GB_CUDA(s, cudaEventElapsedTime(elapsed_ms, s->ev0, s->ev1));
    return GB200_OK;
}

// ---- per-kernel-class timing ----
static cudaEvent_t kt_event(gb200_service* s)
{
    if (!s->kt.pool.empty())
        {
            cudaEvent_t e = s->kt.pool.back();
            s->kt.pool.pop_back();
            return e;
        }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}

void gb_kt_begin(void* ctx, int cls)
{
    gb200_service* s = (gb200_service*)ctx;
    if (!s || !s->kt.enabled) return;
    GbKernelTiming::Pending p;
    p.cls = cls;
    p.a = kt_event(s);
    p.b = kt_event(s);
    cudaEventRecord(p.a, s->stream);
    s->kt.pending.push_back(p);
}

void gb_kt_end(void* ctx)
{
    gb200_service* s = (gb200_service*)ctx;
    if (!s || !s->kt.enabled || s->kt.pending.empty()) return;
    cudaEventRecord(s->kt.pending.back().b, s->stream);
}

static void kt_drain(gb200_service* s)
{
    for (auto& p : s->kt.pending)
        {
            float ms = 0.f;
            if (cudaEventSynchronize(p.b) == cudaSuccess && cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess)
                {
                    s->kt.ms[p.cls] += ms;
                    s->kt.count[p.cls]++;
                }
            s->kt.pool.push_back(p.a);
            s->kt.pool.push_back(p.b);
        }
    s->kt.pending.clear();
}

extern "C" int gb200_service_kernel_timing(gb200_service* s, int enable)
{
    if (!s) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    kt_drain(s);
    s->kt.enabled = enable != 0;
    for (int i = 0; i < GB200_KCLASS_COUNT; i++)
        {
            s->kt.ms[i] = 0;
            s->kt.count[i] = 0;
        }
    return GB200_OK;
}

extern "C" int gb200_service_kernel_stats(gb200_service* s, int kernel_class, double* total_ms, uint64_t* launches)
{
    if (!s || kernel_class < 0 || kernel_class >= GB200_KCLASS_COUNT || !total_ms || !launches) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    kt_drain(s);
    *total_ms = s->kt.ms[kernel_class];
    *launches = s->kt.count[kernel_class];
    return GB200_OK;
}

double gb_fma_peak_launch(float* d_out, int blocks, int iters, cudaStream_t st);

// measured FP32 FMA throughput of this device in TFLOP/s (best of 5 launches, CUDA events)
extern "C" int gb200_service_measure_fp32(gb200_service* s, double* tflops)
{
    if (!s || !tflops) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    const int blocks = s->sm_count * 4;
    void* d;
    int rc = s->scratch(7, sizeof(float) * 512 * (size_t)blocks, &d);
    if (rc) return rc;
    cudaEvent_t a, b;
    GB_CUDA(s, cudaEventCreate(&a));
    GB_CUDA(s, cudaEventCreate(&b));
    double best = 0;
    for (int rep = 0; rep < 6; rep++)
        {
            GB_CUDA(s, cudaEventRecord(a, s->stream));
            const double flops = gb_fma_peak_launch((float*)d, blocks, 4096, s->stream);
            GB_CUDA(s, cudaEventRecord(b, s->stream));
            GB_CUDA(s, cudaEventSynchronize(b));
            float ms = 0;
            GB_CUDA(s, cudaEventElapsedTime(&ms, a, b));
            if (rep > 0 && ms > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
        }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    s->launches += 6;
    *tflops = best;
    return GB200_OK;
}

extern "C" int gb200_service_flush_l2(gb200_service* s, size_t bytes)
{
    if (!s || !bytes) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    if (bytes > s->flush_bytes)
        {
            if (s->d_flush) cudaFree(s->d_flush);
            s->d_flush = nullptr;
            s->flush_bytes = 0;
            GB_CUDA(s, cudaMalloc(&s->d_flush, bytes));
            s->flush_bytes = bytes;
        }
    GB_CUDA(s, cudaMemsetAsync(s->d_flush, 0x5a, bytes, s->stream));
    return GB200_OK;
}

// Tuning switches (profiling / experiments; every default is the measured best).  Keys: "acq_lanes" (1..4), "acq_l2_mb",
// "acq_chunk_t", "acq_fused", "acq_fused_lag", "acq_fused_ring", "acq_fwd_groups" (1..8), "host_threads", "trk_chunk_max", "trace"; and, only before the
// first acquisition handle of a given FFT length exists (they shape the cached plan): "acq_split_n1", "acq_variant", "acq_generic", "acq_fwd_generic".
extern "C" int gb200_service_set_tuning(gb200_service* s, const char* key, int value)
{
    if (!s || !key) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    const std::string k(key);
    GbTuning& t = s->tune;
    if (k == "acq_lanes") t.lanes = value;
    else if (k == "acq_l2_mb") t.l2_mb = value;
    else if (k == "acq_chunk_t") t.chunk_t = value;
    else if (k == "acq_fused") t.fused = value;
    else if (k == "acq_fused_lag") t.fused_lag = value;
    else if (k == "acq_fused_ring") t.fused_ring = value;
    else if (k == "acq_fwd_groups") t.fwd_groups = value;
    else if (k == "host_threads") t.host_threads = value;
    else if (k == "trk_chunk_max") t.trk_chunk_max = value;
    else if (k == "trace") t.trace = value;
    else if (k == "acq_split_n1" || k == "acq_variant" || k == "acq_generic" || k == "acq_fwd_generic")
        {
            if (!s->plans.empty())
                {
                    s->set_error("gb200_service_set_tuning: " + k + " must be set before the first acquisition handle is created");
                    return GB200_E_STATE;
                }
            (k == "acq_split_n1" ? t.split_n1 : (k == "acq_variant" ? t.variant : (k == "acq_generic" ? t.generic : t.fwd_generic))) = value;
        }
    else
        {
            s->set_error("gb200_service_set_tuning: unknown key " + k);
            return GB200_E_INVALID;
        }
    s->acq_ctx.sig.clear();
    return GB200_OK;
}

// ------------------------------------------------------------------ sample ring
static int ring_push(gb200_service* s, int stream_id, const void* src, size_t n_items, int item_type, uint64_t first, cudaMemcpyKind kind)
{
    if (!s || !src) return GB200_E_INVALID;
    if (stream_id < 0 || stream_id >= (int)s->rings.size() || item_type < 0 || item_type > 2)
        {
            s->set_error("gb200_stream_push: stream_id outside [0, " + std::to_string(s->rings.size()) + ") or unknown item_type");
            return GB200_E_INVALID;
        }
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    GbRing& r = s->rings[stream_id];
    const size_t isz = gb_item_bytes(item_type);
    if (!r.d_buf)
        {
            r.cap = s->ring_capacity;
            r.slack = s->ring_slack;
            r.item_type = item_type;
            GB_CUDA(s, cudaMalloc(&r.d_buf, (r.cap + r.slack + GB200_RING_TAIL_ITEMS) * isz));
            GB_CUDA(s, cudaMemsetAsync((char*)r.d_buf + (r.cap + r.slack) * isz, 0, GB200_RING_TAIL_ITEMS * isz, s->copy_stream));
        }
    if (!r.ev_pushed) GB_CUDA(s, cudaEventCreateWithFlags(&r.ev_pushed, cudaEventDisableTiming));
    if (r.item_type != item_type)
        {
            // the front-end was reconfigured: drop what is held and re-type the ring
            GB_CUDA(s, cudaStreamSynchronize(s->copy_stream));
            GB_CUDA(s, cudaStreamSynchronize(s->stream));
            GB_CUDA(s, cudaFree(r.d_buf));
            r.d_buf = nullptr;
            GB_CUDA(s, cudaMalloc(&r.d_buf, (r.cap + r.slack + GB200_RING_TAIL_ITEMS) * isz));
            GB_CUDA(s, cudaMemsetAsync((char*)r.d_buf + (r.cap + r.slack) * isz, 0, GB200_RING_TAIL_ITEMS * isz, s->copy_stream));
            r.item_type = item_type;
            r.any = false;
            r.begin = r.end = 0;
        }
    if (n_items == 0) return GB200_OK;
    if (n_items > r.cap)
        {
            s->set_error("push larger than ring capacity");
            return GB200_E_RANGE;
        }
    const bool restart = !r.any || first != r.end;
    if (r.decim.on && (item_type != GB200_ITEM_32FC || n_items + r.decim.T + r.decim.D > r.cap))
        {
            s->set_error("gb200_stream_push: the stream's decimator needs GB200_ITEM_32FC items and push + taps + decimation <= ring capacity");
            return item_type != GB200_ITEM_32FC ? GB200_E_UNSUPPORTED : GB200_E_RANGE;
        }
    if (r.any && first != r.end)
        {
            // discontinuity: everything held so far is dropped
            r.begin = first;
            for (auto& m : r.marks) m.end = 0;
        }
    else if (!r.any)
        r.begin = first;
    // kernels already launched may still read the items this push overwrites
    if (s->any_consumed) GB_CUDA(s, cudaStreamWaitEvent(s->copy_stream, s->ev_consumed, 0));
    const char* p = (const char*)src;
    uint64_t idx = first;
    size_t left = n_items;
    while (left)
        {
            const uint64_t pos = idx % r.cap;
            const size_t run = (size_t)std::min<uint64_t>(left, r.cap - pos);
            GB_CUDA(s, cudaMemcpyAsync((char*)r.d_buf + pos * isz, p, run * isz, kind, s->copy_stream));
            if (r.conj)
                {
                    GB_CUDA(s, gb_conj_items((char*)r.d_buf + pos * isz, run, item_type, s->copy_stream));
                    s->launches++;
                }
            if (pos < r.slack)  // keep the mirror [cap, cap+slack) == [0, slack) so every window is contiguous
                {
                    const size_t mrun = (size_t)std::min<uint64_t>(run, r.slack - pos);
                    GB_CUDA(s, cudaMemcpyAsync((char*)r.d_buf + (r.cap + pos) * isz, (const char*)r.d_buf + pos * isz, mrun * isz, cudaMemcpyDeviceToDevice, s->copy_stream));
                }
            p += run * isz;
            idx += run;
            left -= run;
        }
    GB_CUDA(s, cudaEventRecord(r.ev_pushed, s->copy_stream));
    r.pushed_unseen = true;
    {
        GbRing::Mark& m = r.marks[r.push_seq % GbRing::kMarks];
        if (!m.ev) GB_CUDA(s, cudaEventCreateWithFlags(&m.ev, cudaEventDisableTiming));
        GB_CUDA(s, cudaEventRecord(m.ev, s->copy_stream));
        m.end = first + n_items;
        m.seq = ++r.push_seq;
    }
    r.end = first + n_items;
    r.any = true;
    if (r.end - r.begin > r.cap) r.begin = r.end - r.cap;
    if (r.decim.on)
        {
            // fir_filter_ccf(decimation, taps) of the flowgraph's acquisition resampler, behind the copy on the same stream
            GbRing::Decimator& d = r.decim;
            if (restart)
                {
                    d.valid_begin = first;
                    d.next_k = (first + d.D - 1) / d.D;
                }
            const uint64_t newest = r.end - 1;
            if (d.next_k * d.D <= newest)
                {
                    const uint64_t k_last = newest / d.D;
                    const unsigned n_out = (unsigned)(k_last - d.next_k + 1);
                    int l = 0;
                    GB_CUDA(s, gb_fir_decimate((const float2*)r.d_buf, r.cap, d.valid_begin, d.next_k, n_out, d.D, d.d_taps, d.T, d.d_out, s->copy_stream, &l));
                    s->launches += l;
                    const uint64_t k_first = d.next_k;
                    d.next_k = k_last + 1;
                    const int dst = d.dst;
                    const float2* src_out = d.d_out;
                    // (r may dangle after this call if the rings vector were resized; it is not: fixed size from creation)
                    return ring_push(s, dst, src_out, n_out, GB200_ITEM_32FC, k_first, cudaMemcpyDeviceToDevice);
                }
        }
    return GB200_OK;
}

extern "C" int gb200_stream_set_decimator(gb200_service* s, int src_stream, int dst_stream, uint32_t decimation, const float* taps, uint32_t n_taps)
{
    if (!s) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    if (src_stream < 0 || src_stream >= (int)s->rings.size())
        {
            s->set_error("gb200_stream_set_decimator: src_stream out of range");
            return GB200_E_INVALID;
        }
    GbRing::Decimator& d = s->rings[src_stream].decim;
    GB_CUDA(s, cudaStreamSynchronize(s->copy_stream));
    if (d.d_taps) cudaFree(d.d_taps);
    if (d.d_out) cudaFree(d.d_out);
    d = GbRing::Decimator();
    if (decimation == 0) return GB200_OK;
    if (dst_stream < 0 || dst_stream >= (int)s->rings.size() || dst_stream == src_stream || !taps || n_taps == 0 || s->rings[dst_stream].decim.on)
        {
            s->set_error("gb200_stream_set_decimator: dst_stream must be another stream of the service without a decimator of its own, taps non-empty");
            return GB200_E_INVALID;
        }
    const size_t span_min = sizeof(float2) * (size_t)decimation * (((size_t)31 * decimation + n_taps + decimation - 1) / decimation);
    if (span_min > 96 * 1024 || (uint64_t)n_taps + decimation >= s->ring_capacity)
        {
            s->set_error("gb200_stream_set_decimator: taps x decimation too large for one CTA's staging buffer (96 KB) or the ring");
            return GB200_E_UNSUPPORTED;
        }
    d.out_cap = (size_t)(s->ring_capacity / decimation) + 2;
    GB_CUDA(s, cudaMalloc(&d.d_taps, sizeof(float) * n_taps));
    GB_CUDA(s, cudaMalloc(&d.d_out, sizeof(float2) * d.out_cap));
    GB_CUDA(s, cudaMemcpy(d.d_taps, taps, sizeof(float) * n_taps, cudaMemcpyHostToDevice));
    d.D = decimation;
    d.T = n_taps;
    d.dst = dst_stream;
    d.next_k = 0;
    d.valid_begin = 0;
    d.on = true;
    // the next push starts a fresh filter run whatever the stream held before
    GbRing& r = s->rings[src_stream];
    if (r.any)
        {
            d.valid_begin = r.end;
            d.next_k = (r.end + d.D - 1) / d.D;
        }
    return GB200_OK;
}

// = gr::filter::firdes::low_pass(gain, fs, fc, tw, WIN_HAMMING) (GNU Radio >= 3.7.3; not in the reference tree): windowed sinc,
// tap count from the window's 53 dB attenuation, Hamming window gr::fft::window::hamming, taps scaled to DC gain `gain`.
extern "C" int gb200_firdes_low_pass(double gain, double sampling_freq, double cutoff_freq, double transition_width, float* taps, uint32_t* n_taps)
{
    if (!taps || !n_taps || !(sampling_freq > 0.0) || !(cutoff_freq > 0.0) || cutoff_freq > sampling_freq / 2 || !(transition_width > 0.0)) return GB200_E_INVALID;
    int ntaps = (int)(53.0 * sampling_freq / (22.0 * transition_width));
    if ((ntaps & 1) == 0) ntaps++;
    if ((uint32_t)ntaps > *n_taps)
        {
            *n_taps = (uint32_t)ntaps;
            return GB200_E_RANGE;
        }
    std::vector<float> w(ntaps);
    {
        const float M = static_cast<float>(ntaps - 1);
        for (int n = 0; n < ntaps; n++) w[n] = 0.54 - 0.46 * cos((2 * M_PI * n) / M);
    }
    const int M = (ntaps - 1) / 2;
    const double fwT0 = 2 * M_PI * cutoff_freq / sampling_freq;
    for (int n = -M; n <= M; n++)
        {
            if (n == 0)
                taps[n + M] = fwT0 / M_PI * w[n + M];
            else
                taps[n + M] = sin(n * fwT0) / (n * M_PI) * w[n + M];
        }
    double fmax = taps[0 + M];
    for (int n = 1; n <= M; n++) fmax += 2 * taps[n + M];
    gain /= fmax;
    for (int i = 0; i < ntaps; i++) taps[i] *= gain;
    *n_taps = (uint32_t)ntaps;
    return GB200_OK;
}

extern "C" int gb200_stream_read(gb200_service* s, int stream_id, uint64_t first_sample_index, size_t n_items, void* dst)
{
    if (!s || !dst) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    if (stream_id < 0 || stream_id >= (int)s->rings.size() || !s->rings[stream_id].d_buf)
        {
            s->set_error("gb200_stream_read: no such stream");
            return GB200_E_INVALID;
        }
    GbRing& r = s->rings[stream_id];
    const size_t isz = gb_item_bytes(r.item_type);
    if (!r.any || first_sample_index < r.begin || first_sample_index + n_items > r.end)
        {
            s->set_error("gb200_stream_read: items not held by the ring");
            return GB200_E_RANGE;
        }
    GB_CUDA(s, cudaStreamSynchronize(s->copy_stream));
    uint64_t idx = first_sample_index;
    size_t left = n_items;
    char* p = (char*)dst;
    while (left)
        {
            const uint64_t pos = idx % r.cap;
            const size_t run = (size_t)std::min<uint64_t>(left, r.cap - pos);
            GB_CUDA(s, cudaMemcpy(p, (const char*)r.d_buf + pos * isz, run * isz, cudaMemcpyDeviceToHost));
            p += run * isz;
            idx += run;
            left -= run;
        }
    return GB200_OK;
}

extern "C" int gb200_stream_set_inverted_spectrum(gb200_service* s, int stream_id, int inverted)
{
    if (!s || stream_id < 0 || stream_id >= (int)s->rings.size()) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    s->rings[stream_id].conj = inverted != 0;
    return GB200_OK;
}

extern "C" int gb200_stream_push(gb200_service* s, int stream_id, const void* samples, size_t n_items, int item_type, uint64_t first_sample_index)
{
    return ring_push(s, stream_id, samples, n_items, item_type, first_sample_index, cudaMemcpyHostToDevice);
}

// Several blocks of one flowgraph are fed the SAME conditioned stream by the scheduler (every channel's acquisition and tracking
// block hangs off the signal conditioner's output): each hands over what it was given, the first to arrive copies it to the
// device, the others find their items already in the ring.
extern "C" int gb200_stream_push_shared(gb200_service* s, int stream_id, const void* samples, size_t n_items, int item_type, uint64_t first_sample_index)
{
    if (!s || !samples) return GB200_E_INVALID;
    if (stream_id < 0 || stream_id >= (int)s->rings.size() || item_type < 0 || item_type > 2) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    const GbRing& r = s->rings[stream_id];
    if (r.any && r.item_type == item_type && first_sample_index <= r.end)
        {
            const uint64_t end = first_sample_index + n_items;
            if (end <= r.end) return GB200_OK;  // all of it is there already (or older than what the ring still holds)
            const uint64_t skip = r.end - first_sample_index;
            return ring_push(s, stream_id, (const char*)samples + skip * gb_item_bytes(item_type), (size_t)(n_items - skip), item_type, r.end, cudaMemcpyHostToDevice);
        }
    return ring_push(s, stream_id, samples, n_items, item_type, first_sample_index, cudaMemcpyHostToDevice);
}

extern "C" int gb200_stream_wait(gb200_service* s, int stream_id)
{
    if (!s || stream_id < 0 || stream_id >= (int)s->rings.size()) return GB200_E_INVALID;
    cudaEvent_t ev;
    {
        std::lock_guard<std::recursive_mutex> lk(s->mu);
        ev = s->rings[stream_id].ev_pushed;
    }
    if (!ev) return GB200_OK;  // nothing pushed yet
    cudaSetDevice(s->device);
    GB_CUDA(s, cudaEventSynchronize(ev));
    return GB200_OK;
}

extern "C" int gb200_stream_push_device(gb200_service* s, int stream_id, const void* d_samples, size_t n_items, int item_type, uint64_t first_sample_index)
{
    return ring_push(s, stream_id, d_samples, n_items, item_type, first_sample_index, cudaMemcpyDeviceToDevice);
}

// device pointer to `len` contiguous items starting at absolute index `first`, or GB200_E_RANGE
static int ring_window(gb200_service* s, int stream_id, uint64_t first, uint64_t len, int item_type, const void** out)
{
    if (stream_id < 0 || stream_id >= (int)s->rings.size()) return GB200_E_INVALID;
    GbRing& r = s->rings[stream_id];
    if (!r.any || first < r.begin || first + len > r.end || len > r.slack)
        {
            char msg[200];
            snprintf(msg, sizeof msg, "samples [%llu, %llu) not in ring [%llu, %llu) (slack %llu)", (unsigned long long)first,
                (unsigned long long)(first + len), (unsigned long long)r.begin, (unsigned long long)r.end, (unsigned long long)r.slack);
            s->set_error(msg);
            return GB200_E_RANGE;
        }
    if (r.item_type != item_type)
        {
            s->set_error("handle item_type differs from the stream's");
            return GB200_E_INVALID;
        }
    *out = (const char*)r.d_buf + (first % r.cap) * gb_item_bytes(item_type);
    return GB200_OK;
}

// ------------------------------------------------------------------ FFT plans
static bool factor_radices(int M, int* rad, int* nrad)
{
    int n = 0, m = M;
    const int pref[] = {8, 4, 5, 3, 2};
    for (int p : pref)
        while (m % p == 0 && m > 1)
            {
                if (n >= GB200_MAX_RADIX_PASSES) return false;
                rad[n++] = p;
                m /= p;
            }
    for (int p = 7; m > 1; p += 2)
        {
            while (m % p == 0)
                {
                    if (p > GB200_MAX_GENERIC_RADIX || n >= GB200_MAX_RADIX_PASSES) return false;
                    rad[n++] = p;
                    m /= p;
                }
            if (p > GB200_MAX_GENERIC_RADIX && m > 1) return false;
        }
    if (M == 1) n = 0;
    *nrad = n;
    return true;
}

static int pick_tile(int M)
{
    // shared memory per CTA = (2*M*(tile+1) + 2*M) * 8 B; keep <= ~72 KB when possible (3 CTAs/SM), hard cap 200 KB
    const int tiles[] = {16, 8, 4, 2, 1};
    for (int t : tiles)
        if (((size_t)2 * M * (t + 1) + 2 * (size_t)M) * 8 <= 72 * 1024) return t;
    for (int t : tiles)
        if (((size_t)2 * M * (t + 1) + 2 * (size_t)M) * 8 <= 200 * 1024) return t;
    return 0;
}

static int get_plan(gb200_service* s, int N, const GbFftPlan** out)
{
    auto it = s->plans.find(N);
    if (it != s->plans.end())
        {
            *out = &it->second.plan;
            return GB200_OK;
        }
    if (N < 1) return GB200_E_INVALID;
    // N1 <= N2, as close to sqrt(N) as the factorisation allows, both sub-lengths supported
    int best1 = 0;
    for (int n1 = 1; (long long)n1 * n1 <= N; n1++)
        {
            if (N % n1) continue;
            int n2 = N / n1, r[GB200_MAX_RADIX_PASSES], nr;
            if (!factor_radices(n1, r, &nr) || !factor_radices(n2, r, &nr)) continue;
            if (!pick_tile(n1) || !pick_tile(n2)) continue;
            best1 = n1;
        }
    {
        // prefer a split with compile-time inverse kernels (tuning key "acq_split_n1" steers away from it)
        int p1 = 0, p2 = 0, r[GB200_MAX_RADIX_PASSES], nr;
        if (!s->tune.generic)
            {
                bool ok = false;
                if (s->tune.split_n1 > 0 && N % s->tune.split_n1 == 0)
                    {
                        p1 = s->tune.split_n1;
                        p2 = N / p1;
                        ok = true;
                    }
                else
                    ok = gb_acq_preferred_split(N, 0, &p1, &p2);
                if (ok && factor_radices(p1, r, &nr) && factor_radices(p2, r, &nr) && pick_tile(p1) && pick_tile(p2)) best1 = p1;
            }
    }
    if (!best1)
        {
            s->set_error("FFT length " + std::to_string(N) + " unsupported (prime factor > 64 or sub-transform too long)");
            return GB200_E_UNSUPPORTED;
        }
    GbPlanEntry e;
    memset(&e.plan, 0, sizeof(e.plan));
    e.plan.N = N;
    e.plan.N1 = best1;
    e.plan.N2 = N / best1;
    factor_radices(e.plan.N1, e.plan.rad1, &e.plan.nrad1);
    factor_radices(e.plan.N2, e.plan.rad2, &e.plan.nrad2);
    e.plan.tile1 = pick_tile(e.plan.N1);
    e.plan.tile2 = pick_tile(e.plan.N2);
    e.plan.inv_tile1 = e.plan.tile1;
    {
        // compile-time specialised inverse kernels for the common (N1, N2) pairs; tuning key "acq_generic" forces the general path
        int t1 = 0, t2 = 0;
        int variant = s->tune.variant;
        if (!gb_acq_inverse_tiles(e.plan.N1, e.plan.N2, variant, &t1, &t2)) variant = 0;
        if (!s->tune.generic && gb_acq_inverse_tiles(e.plan.N1, e.plan.N2, variant, &t1, &t2))
            {
                e.plan.inv_fast = 1;
                e.plan.inv_variant = variant;
                e.plan.inv_tile1 = t1;
            }
        e.plan.fwd_generic = s->tune.fwd_generic ? 1 : 0;
    }
    for (int which = 0; which < 4; which++)
        {
            // which 0/1: W_N1^i / W_N2^i (sub-transform twiddles); 2: W_N^i, i < max(256, 16*max(N1,N2)); 3: W_N^(256 i)
            const int M = which == 3 ? (N + 255) / 256 : (which == 2 ? std::max(256, 16 * std::max(e.plan.N1, e.plan.N2)) : (which ? e.plan.N2 : e.plan.N1));
            const int D = which >= 2 ? N : M;
            const long long mul = which == 3 ? 256 : 1;
            std::vector<float2> tw(M);
            for (int i = 0; i < M; i++)
                {
                    const double a = -6.283185307179586476925286766559 * (double)((i * mul) % D) / (double)D;
                    tw[i] = make_float2((float)cos(a), (float)sin(a));
                }
            float2* d = nullptr;
            GB_CUDA(s, cudaMalloc(&d, sizeof(float2) * M));
            GB_CUDA(s, cudaMemcpyAsync(d, tw.data(), sizeof(float2) * M, cudaMemcpyHostToDevice, s->stream));
            GB_CUDA(s, cudaStreamSynchronize(s->stream));
            if (which == 3)
                e.d_twNhi = d;
            else if (which == 2)
                e.d_twN = d;
            else if (which)
                e.d_tw2 = d;
            else
                e.d_tw1 = d;
        }
    e.plan.tw1 = e.d_tw1;
    e.plan.tw2 = e.d_tw2;
    e.plan.twN = e.d_twN;
    e.plan.twNhi = e.d_twNhi;
    auto ins = s->plans.emplace(N, e);
    *out = &ins.first->second.plan;
    return GB200_OK;
}

// ------------------------------------------------------------------ acquisition
struct gb200_acq
{
    gb200_service* s = nullptr;
    gb200_acq_cfg c;
    uint32_t consumed = 0, fft_size = 0, num_bins = 0, eff = 0;
    bool use_cfar = false;
    bool keep_grid = false;
    const GbFftPlan* plan = nullptr;
    float2* d_code_fft = nullptr;
    bool code_set = false;
    float* d_grid = nullptr;  // [max(num_bins, bins_step2)][eff]
    // End of an attempt (pcps_acquisition.cc:912-929) zeroes the grid.  The next first dwell overwrites every row it searches, so the
    // zeroing is only owed to whoever looks at the grid before that: it is recorded here and done by grid_settle() on demand.
    bool grid_zero_pending = false;
    size_t grid_rows = 0;
    float* d_tmp = nullptr;   // d_tmp_buffer image, fft_size floats (non-CFAR statistic only)
    // block state (pcps_acquisition members)
    uint32_t dwell_counter = 0;  // d_num_noncoherent_integrations_counter
    bool step_two = false;
    int positive_acq = 0;
    float doppler_center_step_two = 0.f;
    // wipe-off description cache (recomputed when the Doppler grid changes)
    std::vector<float> seg_steps;
    std::vector<GbWipeSeg> segs;
    std::vector<int> seg_offs;
    // launch -> collect hand-over
    uint64_t pending_first = 0;
    int pending_group = -1, pending_slot = -1;
    bool pending = false;
    uint64_t cfg_epoch = 0;  // bumped by every setter that changes grouping, tables or device pointers (batch signature)

    // sibling blocks served through the same Doppler loop
    std::vector<float> custom_steps;  // gb200_acq_set_doppler_grid: phase step per bin replaces the doppler_max/doppler_step rule
    uint32_t fold = 1;                // gb200_acq_create_folded: QuickSync, consumed = fold * fft_size
    int scaled_accum = 0;             // gb200_acq_set_scaled_accumulation: 1 = Tong detector, grid += mag / (N^4 input_power) every dwell; 2 = plain grid += mag

    bool need_grid() const { return keep_grid || !use_cfar || c.max_dwells > 1 || scaled_accum; }
};

static const double kGpsTwoPi = 6.283185307179586;  // GPS_L1_CA.h:46

// pcps_acquisition.cc:296-305: phase_step_rad as float32
static float acq_phase_step(const gb200_acq* a, float freq) { return (float)(kGpsTwoPi * freq / (float)a->c.fs_in); }

static void acq_bin_steps(const gb200_acq* a, std::vector<float>& steps)
{
    steps.clear();
    if (!a->step_two && !a->custom_steps.empty())
        steps = a->custom_steps;
    else if (!a->step_two)
        {
            for (uint32_t i = 0; i < a->num_bins; i++)
                {
                    const int32_t doppler = -(int32_t)a->c.doppler_max + (int32_t)(a->c.doppler_step * i);  // :357-358
                    steps.push_back(acq_phase_step(a, (float)(a->c.freq_offset_hz + doppler)));
                }
        }
    else
        {
            for (uint32_t i = 0; i < a->c.num_doppler_bins_step2; i++)  // :383-390
                {
                    const float doppler = ((float)i - (float)floor(a->c.num_doppler_bins_step2 / 2.0)) * a->c.doppler_step2;
                    steps.push_back(acq_phase_step(a, a->doppler_center_step_two + doppler));
                }
        }
}

// Simulate volk_gnsssdr_s32f_sincos_32fc_generic's float32 phase accumulation (phase starts at 0, increment
// -phase_step_rad, pcps_acquisition.cc:307-309) and compress it into exact piecewise-linear segments.
static void acq_wipe_segments(float phase_step_rad, uint32_t n, std::vector<GbWipeSeg>& out)
{
    const float inc = -phase_step_rad;
    volatile float ph = 0.0f;  // volatile: keep every partial sum rounded to float32
    GbWipeSeg cur;
    cur.n_start = 0;
    cur.ph_start = 0.0f;
    cur.inc_eff = 0.0f;
    bool have_inc = false;
    for (uint32_t i = 0; i + 1 < n; i++)
        {
            const float prev = ph;
            ph = prev + inc;
            const float next = ph;
            const float d = (float)((double)next - (double)prev);
            if (!have_inc)
                {
                    cur.inc_eff = d;
                    have_inc = true;
                }
            else if (d != cur.inc_eff)
                {
                    out.push_back(cur);
                    cur.n_start = (int)i;
                    cur.ph_start = prev;
                    cur.inc_eff = d;
                }
        }
    out.push_back(cur);
}

static void acq_all_segments(const std::vector<float>& steps, uint32_t n, std::vector<GbWipeSeg>& segs, std::vector<int>& offs)
{
    segs.clear();
    offs.clear();
    for (float st : steps)
        {
            offs.push_back((int)segs.size());
            acq_wipe_segments(st, n, segs);
        }
    offs.push_back((int)segs.size());
}

static void acq_derive(gb200_acq* a)
{
    const gb200_acq_cfg& c = a->c;
    a->consumed = (uint32_t)((float)c.sampled_ms * c.samples_per_ms * (float)(c.bit_transition_flag ? 2 : 1));  // :75
    a->fft_size = (c.sampled_ms == c.ms_per_code) ? a->consumed : a->consumed * 2;                               // :76-84
    if (c.bit_transition_flag)
        {
            a->fft_size = a->consumed * 2;  // :113-117
            a->c.max_dwells = 1;
        }
    a->eff = c.bit_transition_flag ? a->fft_size / 2 : a->fft_size;
    a->use_cfar = (a->c.max_dwells == 1) ? (c.use_CFAR_algorithm_flag != 0) : false;  // :152-159
    a->num_bins = c.doppler_step ? (uint32_t)std::ceil((double)((int32_t)c.doppler_max - (int32_t)(-(int32_t)c.doppler_max)) / (double)c.doppler_step) : 0;  // :326
    a->custom_steps.clear();
}

static int acq_alloc_grid(gb200_acq* a)
{
    gb200_service* s = a->s;
    const size_t rows = std::max<size_t>(a->num_bins, a->c.make_2_steps ? a->c.num_doppler_bins_step2 : 0);
    if (a->need_grid() && (a->d_grid == nullptr || rows > a->grid_rows))
        {
            if (a->d_grid) cudaFree(a->d_grid);
            a->d_grid = nullptr;
            GB_CUDA(s, cudaMalloc(&a->d_grid, sizeof(float) * rows * a->eff));
            GB_CUDA(s, cudaMemsetAsync(a->d_grid, 0, sizeof(float) * rows * a->eff, s->stream));
            a->grid_rows = rows;
        }
    if (!a->use_cfar && !a->d_tmp)
        {
            GB_CUDA(s, cudaMalloc(&a->d_tmp, sizeof(float) * a->fft_size));
            GB_CUDA(s, cudaMemsetAsync(a->d_tmp, 0, sizeof(float) * a->fft_size, s->stream));
        }
    return GB200_OK;
}

extern "C" int gb200_acq_create(gb200_service* s, const gb200_acq_cfg* cfg, gb200_acq** out) { return gb200_acq_create_folded(s, cfg, 1, out); }

extern "C" int gb200_acq_create_folded(gb200_service* s, const gb200_acq_cfg* cfg, uint32_t fold, gb200_acq** out)
{
    if (!s || !cfg || !out || fold == 0) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    *out = nullptr;
    if (cfg->item_type < 0 || cfg->item_type > 2 || cfg->doppler_step == 0 || cfg->fs_in <= 0 || cfg->max_dwells == 0)
        {
            s->set_error("gb200_acq_create: bad configuration");
            return GB200_E_INVALID;
        }
    gb200_acq* a = new gb200_acq();
    a->s = s;
    a->c = *cfg;
    acq_derive(a);
    if (fold > 1)
        {
            // pcps_quicksync_acquisition_cc.cc:94-95: d_fft_size = samples_per_code / folding_factor, fed by folding_factor^2 segments
            if (cfg->bit_transition_flag || cfg->make_2_steps || cfg->sampled_ms != cfg->ms_per_code || a->consumed % fold != 0)
                {
                    s->set_error("gb200_acq_create_folded: needs sampled_ms == ms_per_code, no bit transition / two steps, and consumed samples divisible by fold");
                    delete a;
                    return GB200_E_INVALID;
                }
            a->fold = fold;
            a->fft_size = a->consumed / fold;
            a->eff = a->fft_size;
        }
    if (a->consumed == 0 || a->num_bins == 0 || (uint64_t)a->consumed > s->ring_slack)
        {
            s->set_error("gb200_acq_create: consumed samples is 0 or exceeds the ring slack");
            delete a;
            return GB200_E_INVALID;
        }
    int rc = get_plan(s, (int)a->fft_size, &a->plan);
    if (rc != GB200_OK)
        {
            delete a;
            return rc;
        }
    cudaError_t e = cudaMalloc(&a->d_code_fft, sizeof(float2) * a->fft_size);
    if (e != cudaSuccess)
        {
            s->set_error(std::string("cudaMalloc code fft: ") + cudaGetErrorString(e));
            delete a;
            return GB200_E_CUDA;
        }
    rc = acq_alloc_grid(a);
    if (rc != GB200_OK)
        {
            gb200_acq_destroy(a);
            return rc;
        }
    *out = a;
    return GB200_OK;
}

extern "C" int gb200_acq_destroy(gb200_acq* a)
{
    if (!a) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(a->s->mu);
    cudaSetDevice(a->s->device);
    cudaStreamSynchronize(a->s->stream);
    a->s->acq_ctx.sig.clear();  // a later handle may reuse this address
    if (a->d_code_fft) cudaFree(a->d_code_fft);
    if (a->d_grid) cudaFree(a->d_grid);
    if (a->d_tmp) cudaFree(a->d_tmp);
    delete a;
    return GB200_OK;
}

extern "C" uint32_t gb200_acq_fft_size(const gb200_acq* a) { return a ? a->fft_size : 0; }
extern "C" uint32_t gb200_acq_consumed_samples(const gb200_acq* a) { return a ? a->consumed : 0; }
extern "C" uint32_t gb200_acq_num_doppler_bins(const gb200_acq* a) { return a ? a->num_bins : 0; }

extern "C" int gb200_acq_set_local_code(gb200_acq* a, const float* code_iq, size_t n_items)
{
    if (!a || !code_iq) return GB200_E_INVALID;
    gb200_service* s = a->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    const uint32_t N = a->fft_size;
    // zero-pad layout of pcps_acquisition.cc:253-270
    size_t offset, count;
    if (a->c.bit_transition_flag)
        {
            offset = N / 2;
            count = N / 2;
        }
    else if (a->fold > 1)
        {
            offset = 0;
            count = N;  // the caller's folded code (pcps_quicksync_acquisition_cc.cc:186-196)
        }
    else if (a->c.sampled_ms == a->c.ms_per_code)
        {
            offset = 0;
            count = a->consumed;
        }
    else
        {
            offset = N - a->consumed;
            count = a->consumed;
        }
    if (n_items < count)
        {
            s->set_error("gb200_acq_set_local_code: code shorter than the block's copy length");
            return GB200_E_INVALID;
        }
    void* hp;
    int rc = s->pinned(sizeof(float2) * N, &hp);
    if (rc) return rc;
    GB_CUDA(s, cudaStreamSynchronize(s->stream));  // the pinned staging buffer may still be in flight
    memset(hp, 0, sizeof(float2) * N);
    memcpy((float2*)hp + offset, code_iq, sizeof(float2) * count);
    void *d_in, *d_col;
    if ((rc = s->scratch(2, sizeof(float2) * N, &d_in))) return rc;
    if ((rc = s->scratch(3, sizeof(float2) * N, &d_col))) return rc;
    GB_CUDA(s, cudaMemcpyAsync(d_in, hp, sizeof(float2) * N, cudaMemcpyHostToDevice, s->stream));
    GbAcqFwdParams p;
    p.samples = d_in;
    p.consumed = (int)N;
    p.item_type = GB200_ITEM_32FC;
    p.segs = nullptr;
    p.seg_offsets = nullptr;
    p.colbuf = (float2*)d_col;
    p.spectra = a->d_code_fft;
    int l = 0;
    GB_CUDA(s, gb_acq_forward(*a->plan, p, 1, false, s->stream, &l, nullptr));
    GB_CUDA(s, gb_acq_conj(a->d_code_fft, N, s->stream, &l));  // volk_32fc_conjugate_32fc, :273
    s->launches += l;
    GB_CUDA(s, cudaStreamSynchronize(s->stream));
    a->code_set = true;
    return GB200_OK;
}

extern "C" int gb200_acq_set_doppler(gb200_acq* a, uint32_t doppler_max, uint32_t doppler_step, int64_t freq_offset_hz)
{
    if (!a || doppler_step == 0) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(a->s->mu);
    cudaSetDevice(a->s->device);
    a->c.doppler_max = doppler_max;
    a->c.doppler_step = doppler_step;
    a->c.freq_offset_hz = freq_offset_hz;
    a->cfg_epoch++;
    acq_derive(a);
    return acq_alloc_grid(a);
}

extern "C" int gb200_acq_set_doppler_grid(gb200_acq* a, uint32_t n_bins, const float* phase_step_rad)
{
    if (!a || !phase_step_rad || n_bins == 0 || n_bins > 65535) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(a->s->mu);
    cudaSetDevice(a->s->device);
    if (a->c.make_2_steps)
        {
            a->s->set_error("gb200_acq_set_doppler_grid: not combined with make_2_steps");
            return GB200_E_INVALID;
        }
    a->custom_steps.assign(phase_step_rad, phase_step_rad + n_bins);
    a->num_bins = n_bins;
    a->cfg_epoch++;
    return acq_alloc_grid(a);
}

extern "C" int gb200_acq_set_scaled_accumulation(gb200_acq* a, int enable)
{
    if (!a) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(a->s->mu);
    cudaSetDevice(a->s->device);
    if (a->c.bit_transition_flag || a->c.make_2_steps)
        {
            a->s->set_error("gb200_acq_set_scaled_accumulation: not combined with bit_transition_flag / make_2_steps");
            return GB200_E_INVALID;
        }
    a->scaled_accum = enable == 2 ? 2 : (enable != 0 ? 1 : 0);
    a->cfg_epoch++;
    return acq_alloc_grid(a);
}

extern "C" int gb200_acq_set_threshold(gb200_acq* a, float threshold)
{
    if (!a) return GB200_E_INVALID;
    a->c.threshold = threshold;
    return GB200_OK;
}

extern "C" int gb200_acq_keep_grid(gb200_acq* a, int keep)
{
    if (!a) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(a->s->mu);
    cudaSetDevice(a->s->device);
    a->keep_grid = keep != 0;
    a->cfg_epoch++;
    return acq_alloc_grid(a);
}

static cudaError_t grid_settle(gb200_acq* a)
{
    if (!a->grid_zero_pending || !a->d_grid) return cudaSuccess;
    a->grid_zero_pending = false;
    return cudaMemsetAsync(a->d_grid, 0, sizeof(float) * a->grid_rows * a->eff, a->s->stream);
}

extern "C" int gb200_acq_reset(gb200_acq* a)
{
    if (!a) return GB200_E_INVALID;
    gb200_service* s = a->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    a->dwell_counter = 0;
    a->step_two = false;
    a->positive_acq = 0;
    a->grid_zero_pending = false;
    if (a->d_grid) GB_CUDA(s, cudaMemsetAsync(a->d_grid, 0, sizeof(float) * a->grid_rows * a->eff, s->stream));
    return GB200_OK;
}

extern "C" int gb200_acq_read_grid(gb200_acq* a, float* dst)
{
    if (!a || !dst) return GB200_E_INVALID;
    gb200_service* s = a->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    if (!a->d_grid)
        {
            s->set_error("gb200_acq_read_grid: handle keeps no grid (see gb200_acq_keep_grid)");
            return GB200_E_STATE;
        }
    GB_CUDA(s, grid_settle(a));
    GB_CUDA(s, cudaMemcpyAsync(dst, a->d_grid, sizeof(float) * a->num_bins * a->eff, cudaMemcpyDeviceToHost, s->stream));
    GB_CUDA(s, cudaStreamSynchronize(s->stream));
    return GB200_OK;
}

extern "C" int gb200_acq_iq_combine(gb200_acq* ia, gb200_acq* qa, gb200_acq* ib, gb200_acq* qb, float norm, gb200_acq_iq_bin* out)
{
    if (!ia || !out) return GB200_E_INVALID;
    gb200_service* s = ia->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    gb200_acq* h[4] = {ia, qa, ib, qb};
    const float* rows[4] = {nullptr, nullptr, nullptr, nullptr};
    for (int k = 0; k < 4; k++)
        {
            if (!h[k]) continue;
            if (h[k]->s != s || h[k]->eff != ia->eff || h[k]->num_bins != ia->num_bins)
                {
                    s->set_error("gb200_acq_iq_combine: handles differ in service, geometry or Doppler grid");
                    return GB200_E_INVALID;
                }
            if (!h[k]->d_grid || !h[k]->keep_grid)
                {
                    s->set_error("gb200_acq_iq_combine: every handle must keep its grid (gb200_acq_keep_grid) and have run a dwell");
                    return GB200_E_STATE;
                }
            GB_CUDA(s, grid_settle(h[k]));
            rows[k] = h[k]->d_grid;
        }
    if (qb && !ib)
        {
            s->set_error("gb200_acq_iq_combine: Q_B needs I_B (the block compares magt_QB on the I_B row)");
            return GB200_E_INVALID;
        }
    void* d_out;
    int rc;
    const size_t bytes = sizeof(gb200_acq_iq_bin) * ia->num_bins;
    if ((rc = s->scratch(9, bytes, &d_out))) return rc;
    int l = 0;
    GB_CUDA(s, gb_acq_iq_combine(rows, (int)ia->num_bins, (int)ia->eff, norm, (gb200_acq_iq_bin*)d_out, s->stream, &l));
    s->launches += l;
    GB_CUDA(s, cudaMemcpyAsync(out, d_out, bytes, cudaMemcpyDeviceToHost, s->stream));
    GB_CUDA(s, cudaStreamSynchronize(s->stream));
    return GB200_OK;
}

// pcps_acquisition_fine_doppler_cc::compute_CAF (:296-322): second peak of one kept grid row with the +-half_width exclusion
// around `idx` (same wrap rule and do-while as pcps_acquisition's, but on the whole row)
extern "C" int gb200_acq_second_peak(gb200_acq* a, uint32_t bin, uint32_t idx, uint32_t half_width, float* second_peak)
{
    if (!a || !second_peak) return GB200_E_INVALID;
    gb200_service* s = a->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    if (!a->d_grid || bin >= a->num_bins || idx >= a->eff)
        {
            s->set_error("gb200_acq_second_peak: handle keeps no grid, or bin / index out of range");
            return GB200_E_STATE;
        }
    GB_CUDA(s, grid_settle(a));
    void* d_tmp;
    int rc;
    if ((rc = s->scratch(18, sizeof(float) * (a->eff + 4), &d_tmp))) return rc;
    int l = 0;
    float* d_second = (float*)d_tmp + a->eff;
    GB_CUDA(s, gb_acq_second_peak_row(a->d_grid + (size_t)bin * a->eff, (float*)d_tmp, (int)a->eff, (int)idx, (int)half_width, d_second, s->stream, &l));
    s->launches += l;
    GB_CUDA(s, cudaMemcpyAsync(second_peak, d_second, sizeof(float), cudaMemcpyDeviceToHost, s->stream));
    GB_CUDA(s, cudaStreamSynchronize(s->stream));
    return GB200_OK;
}

// pcps_acquisition_fine_doppler_cc::estimate_Doppler (:400-441): items x replica, zero-padded to fft_len, forward FFT, |.|^2,
// first maximum in natural frequency order
extern "C" int gb200_acq_wiped_spectrum_peak(gb200_service* s, int stream_id, uint64_t first_sample_index, const float* replica_iq, uint32_t n_items,
    uint32_t fft_len, uint32_t* peak_index, float* peak_mag)
{
    if (!s || !replica_iq || !peak_index || !peak_mag || n_items == 0 || fft_len < n_items) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    if (stream_id < 0 || stream_id >= (int)s->rings.size() || !s->rings[stream_id].d_buf)
        {
            s->set_error("gb200_acq_wiped_spectrum_peak: no such stream");
            return GB200_E_INVALID;
        }
    const int item_type = s->rings[stream_id].item_type;
    const GbFftPlan* plan;
    int rc = get_plan(s, (int)fft_len, &plan);
    if (rc) return rc;
    if ((rc = rings_acquire(s))) return rc;
    const void* win;
    if ((rc = ring_window(s, stream_id, first_sample_index, n_items, item_type, &win))) return rc;
    void *d_rp, *d_col, *d_spec, *d_pk;
    if ((rc = s->scratch(18, sizeof(float2) * 2 * (size_t)n_items, &d_rp))) return rc;
    if ((rc = s->scratch(19, sizeof(float2) * (size_t)fft_len, &d_col))) return rc;
    if ((rc = s->scratch(1, sizeof(float2) * (size_t)fft_len, &d_spec))) return rc;
    if ((rc = s->scratch(9, sizeof(GbPeak), &d_pk))) return rc;
    float2* d_rep = (float2*)d_rp;
    float2* d_prod = d_rep + n_items;
    GB_CUDA(s, cudaMemcpyAsync(d_rep, replica_iq, sizeof(float2) * n_items, cudaMemcpyHostToDevice, s->stream));
    int l = 0;
    GB_CUDA(s, gb_acq_code_wipe(win, item_type, d_rep, (int)n_items, d_prod, s->stream, &l));
    GbAcqFwdParams p;
    p.samples = d_prod;
    p.consumed = (int)n_items;
    p.item_type = GB200_ITEM_32FC;
    p.segs = nullptr;
    p.seg_offsets = nullptr;
    p.colbuf = (float2*)d_col;
    p.spectra = (float2*)d_spec;
    GB_CUDA(s, gb_acq_forward(*plan, p, 1, false, s->stream, &l, nullptr));
    GB_CUDA(s, gb_acq_spectrum_peak((const float2*)d_spec, plan->N1, plan->N2, (GbPeak*)d_pk, s->stream, &l));
    s->launches += l;
    GbPeak pk;
    GB_CUDA(s, cudaMemcpyAsync(&pk, d_pk, sizeof(GbPeak), cudaMemcpyDeviceToHost, s->stream));
    GB_CUDA(s, cudaStreamSynchronize(s->stream));
    { int rc0 = rings_release(s); if (rc0) return rc0; }
    *peak_index = pk.idx;
    *peak_mag = pk.mag;
    return GB200_OK;
}

// ---- batched acquisition: launch (asynchronous) / collect ----
namespace
{
size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
}  // namespace

// what decides which column-kernel mode a grid-keeping handle runs in: handles may share a group (one forward side, one launch
// sequence) only when these agree
static int acq_mode_key(const gb200_acq* a)
{
    if (!a->need_grid()) return 0;
    return 1 | (a->use_cfar ? 2 : 0) | (a->scaled_accum << 2) | (a->dwell_counter >= 1 ? 16 : 0) | ((int)a->c.samples_per_chip << 8);
}

extern "C" int gb200_acq_launch_batch(gb200_acq* const* list, int n, int stream_id, uint64_t first_sample_index)
{
    if (!list || n <= 0 || !list[0]) return GB200_E_INVALID;
    gb200_service* s = list[0]->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    AcqBatchCtx& ctx = s->acq_ctx;
    if (ctx.active)
        {
            s->set_error("gb200_acq_launch_batch: previous batch not collected");
            return GB200_E_STATE;
        }
    for (int i = 0; i < n; i++)
        {
            if (!list[i] || list[i]->s != s) return GB200_E_INVALID;
            if (!list[i]->code_set)
                {
                    s->set_error("gb200_acq_run: set_local_code has not been called");
                    return GB200_E_STATE;
                }
            for (int j = 0; j < i; j++)
                if (list[j] == list[i]) return GB200_E_INVALID;
        }
    { int rc0 = rings_acquire(s); if (rc0) return rc0; }
    // Signature of the batch: the handles, what their Doppler grids / modes were last set to, and the state bits that pick a
    // kernel mode.  A launch with the signature of the previous one (the steady state: the same channels searching block after
    // block) reuses the group tables already on the device: no host rebuild, no upload, no wait for the stream.
    std::vector<uint64_t> sig;
    sig.reserve(3 * (size_t)n + 1);
    sig.push_back((uint64_t)n);
    for (int i = 0; i < n; i++)
        {
            sig.push_back((uint64_t)(uintptr_t)list[i]);
            sig.push_back((list[i]->cfg_epoch << 8) | (uint64_t)(list[i]->step_two ? 1 : 0) | ((uint64_t)(acq_mode_key(list[i]) & 0x7f) << 1));
            uint32_t center_bits = 0;  // the step-two grid is centred on the coarse estimate (:383-390)
            if (list[i]->step_two) memcpy(&center_bits, &list[i]->doppler_center_step_two, sizeof center_bits);
            sig.push_back(center_bits);
        }
    const bool reuse = !ctx.sig.empty() && sig == ctx.sig;
    void *h_up = nullptr, *d_up, *d_res;
    int rc;
    if (!reuse)
        {
            ctx.sig.clear();
            // group handles that can share the wiped-off forward spectra
            ctx.groups.clear();
            std::vector<float> steps;
            for (int i = 0; i < n; i++)
                {
                    gb200_acq* a = list[i];
                    acq_bin_steps(a, steps);
                    const int key = acq_mode_key(a);
                    bool placed = false;
                    for (auto& g : ctx.groups)
                        {
                            gb200_acq* b = list[g.members[0]];
                            if (g.mode_key == key && b->fft_size == a->fft_size && b->consumed == a->consumed && b->fold == a->fold && b->c.item_type == a->c.item_type &&
                                b->c.bit_transition_flag == a->c.bit_transition_flag && g.steps == steps)
                                {
                                    g.members.push_back(i);
                                    placed = true;
                                    break;
                                }
                        }
                    if (!placed)
                        {
                            AcqGroup g;
                            g.members.push_back(i);
                            g.steps = steps;
                            g.grid = a->need_grid();
                            g.mode_key = key;
                            ctx.groups.push_back(g);
                        }
                }
            // arenas
            size_t up = 0, res = 0;
            for (auto& g : ctx.groups)
                {
                    gb200_acq* a0 = list[g.members[0]];
                    if (a0->seg_steps != g.steps || a0->seg_offs.empty())
                        {
                            acq_all_segments(g.steps, std::max(a0->fft_size, a0->fold > 1 ? a0->consumed : 0u), a0->segs, a0->seg_offs);
                            a0->seg_steps = g.steps;
                        }
                    g.off_segs = up;
                    up = align_up(up + a0->segs.size() * sizeof(GbWipeSeg), 16);
                    g.off_segoff = up;
                    up = align_up(up + a0->seg_offs.size() * sizeof(int), 16);
                    g.off_code = up;
                    up = align_up(up + g.members.size() * sizeof(void*), 16);
                    g.off_grid = up;
                    up = align_up(up + g.members.size() * sizeof(void*), 16);
                    g.off_tmp = up;
                    up = align_up(up + g.members.size() * sizeof(void*), 16);
                    g.off_power = res;
                    res = align_up(res + 2 * sizeof(float), 16);  // input power, Tong scale
                    g.off_peaks = res;
                    res = align_up(res + g.members.size() * sizeof(GbPeak), 16);
                    g.off_second = res;
                    res = align_up(res + g.members.size() * sizeof(float), 16);
                }
            if ((rc = s->pinned(up, &h_up))) return rc;
            GB_CUDA(s, cudaStreamSynchronize(s->stream));  // the staging buffer / the device tables may still be in use
            if (res > ctx.h_res_bytes)
                {
                    if (ctx.h_res) cudaFreeHost(ctx.h_res);
                    ctx.h_res = nullptr;
                    ctx.h_res_bytes = 0;
                    GB_CUDA(s, cudaMallocHost((void**)&ctx.h_res, 2 * res));
                    ctx.h_res_bytes = 2 * res;
                }
            ctx.res_bytes = res;
            ctx.up_bytes = up;
        }
    if ((rc = s->scratch(4, ctx.up_bytes, &d_up))) return rc;
    if ((rc = s->scratch(5, ctx.res_bytes, &d_res))) return rc;
    if (!reuse)
        {
            for (auto& g : ctx.groups)
                {
                    {
                        gb200_acq* a0 = list[g.members[0]];
                        memcpy((char*)h_up + g.off_segs, a0->segs.data(), a0->segs.size() * sizeof(GbWipeSeg));
                        memcpy((char*)h_up + g.off_segoff, a0->seg_offs.data(), a0->seg_offs.size() * sizeof(int));
                    }
                    for (size_t k = 0; k < g.members.size(); k++)
                        {
                            gb200_acq* a = list[g.members[k]];
                            ((const float2**)((char*)h_up + g.off_code))[k] = a->d_code_fft;
                            ((float**)((char*)h_up + g.off_grid))[k] = g.grid ? a->d_grid : nullptr;
                            ((float**)((char*)h_up + g.off_tmp))[k] = a->d_tmp;
                        }
                }
            GB_CUDA(s, cudaMemcpyAsync(d_up, h_up, ctx.up_bytes, cudaMemcpyHostToDevice, s->stream));
        }

    int l = 0;
    for (auto& g : ctx.groups)
        {
            gb200_acq* a0 = list[g.members[0]];
            const GbFftPlan& plan = *a0->plan;
            const int N = (int)a0->fft_size;
            const int bins = (int)g.steps.size();
            const void* win;
            if ((rc = ring_window(s, stream_id, first_sample_index, a0->consumed, a0->c.item_type, &win))) return rc;
            void *d_col, *d_spec;
            if ((rc = s->scratch(0, sizeof(float2) * (size_t)N * bins, &d_col))) return rc;
            if ((rc = s->scratch(1, sizeof(float2) * (size_t)N * bins, &d_spec))) return rc;
            // input power (CFAR or bit transition, pcps_acquisition.cc:703-709).  It needs only the samples, and only the final
            // statistic needs it: unless it also leaves the |x|^2 image the inverse kernels read, it runs on stream2 beside the
            // forward transforms and is joined before the peak reduction.
            bool want_power = false;
            for (int m : g.members) want_power = want_power || list[m]->use_cfar || list[m]->c.bit_transition_flag || list[m]->scaled_accum;
            float* d_power = (float*)((char*)d_res + g.off_power);
            const bool image = g.grid && !a0->use_cfar;  // the |x|^2 image the block leaves in d_tmp_buffer
            const bool power_aside = want_power && !image && !a0->scaled_accum && !s->kt.enabled;  // (the Tong scale feeds the inverse kernels)
            // folded handles: the power is the mean over all consumed items (pcps_quicksync_acquisition_cc.cc:363-365)
            const int power_n = a0->fold > 1 ? (int)a0->consumed : (int)a0->fft_size;
            if (power_aside)
                {
                    GB_CUDA(s, cudaEventRecord(s->ev_power_fork, s->stream));
                    GB_CUDA(s, cudaStreamWaitEvent(s->stream2, s->ev_power_fork, 0));
                    GB_CUDA(s, gb_acq_input_power(win, (int)a0->consumed, a0->c.item_type, power_n, d_power, nullptr, s->stream2, &l));
                }
            if (want_power && !power_aside)
                {
                    GB_CUDA(s, gb_acq_input_power(win, (int)a0->consumed, a0->c.item_type, power_n, d_power, image ? a0->d_tmp : nullptr, s->stream, &l));
                    if (image)
                        for (size_t k = 1; k < g.members.size(); k++)
                            GB_CUDA(s, cudaMemcpyAsync(list[g.members[k]]->d_tmp, a0->d_tmp, sizeof(float) * N, cudaMemcpyDeviceToDevice, s->stream));
                }
            // a grid whose end-of-attempt zeroing is still owed: a first dwell overwrites all the rows it searches, anything else
            // (accumulation, rows beyond this search) gets the zeroing now
            if (g.grid)
                for (int m : g.members)
                    {
                        gb200_acq* a = list[m];
                        if (!a->grid_zero_pending) continue;
                        if (a->dwell_counter >= 1 || a->grid_rows > (size_t)bins)
                            GB_CUDA(s, grid_settle(a));
                        else
                            a->grid_zero_pending = false;
                    }
            // inverse side in L2-sized chunks of transforms
            const int T = bins * (int)g.members.size();
            // Two lanes: chunk c runs on stream (c & 1) with its own row buffer, so one lane's column kernel overlaps the
            // other lane's row kernel and each kernel's ragged last wave is filled by the other lane.  Both row buffers
            // together stay within the L2 budget.  With per-launch event timing on, everything stays on one stream.
            const int lanes_cfg = std::max(1, std::min(4, s->tune.lanes));
            const size_t l2_total = (size_t)std::max(4, s->tune.l2_mb) << 20;
            const int lanes = s->kt.enabled ? 1 : lanes_cfg;
            const size_t l2_budget = l2_total / (size_t)lanes;
            int chunkT = (int)std::max<size_t>(1, l2_budget / (sizeof(float2) * (size_t)N));
            if (s->tune.chunk_t > 0) chunkT = s->tune.chunk_t;
            chunkT = std::min(chunkT, T);
            const int tiles = (plan.N2 + plan.inv_tile1 - 1) / plan.inv_tile1;
            void* d_rows[4] = {nullptr, nullptr, nullptr, nullptr};
            void* d_part;
            static const int row_slots[4] = {2, 8, 14, 15};
            // lane 0's buffer is sized for the one-lane chunk so that toggling event timing never reallocates
            if ((rc = s->scratch(2, sizeof(float2) * (size_t)N * std::min<size_t>(T, std::max<size_t>(1, l2_total / (sizeof(float2) * (size_t)N))), &d_rows[0]))) return rc;
            for (int k = 1; k < lanes; k++)
                if ((rc = s->scratch(row_slots[k], sizeof(float2) * (size_t)N * chunkT, &d_rows[k]))) return rc;
            if ((rc = s->scratch(3, sizeof(GbPeak) * (size_t)T * tiles, &d_part))) return rc;
            // Opt-in (tuning key "acq_fused"): one persistent launch for the whole inverse side when there is a kernel for this plan
            // and mode (plain arg-max search); otherwise, and by default, chunks of transforms on the lanes below.
            const int fused_cfg = s->tune.fused, fused_lag = std::max(1, s->tune.fused_lag), fused_ring = std::max(2, s->tune.fused_ring);
            GbAcqFwdParams fp;
            fp.samples = win;
            fp.consumed = (int)a0->consumed;
            fp.item_type = a0->c.item_type;
            fp.segs = (const GbWipeSeg*)((char*)d_up + g.off_segs);
            fp.seg_offsets = (const int*)((char*)d_up + g.off_segoff);
            fp.colbuf = (float2*)d_col;
            fp.spectra = (float2*)d_spec;
            fp.fold = (int)a0->fold;
            // Forward side: one launch pair on the main stream.  Tuning key "acq_fwd_groups" > 1: in that many groups of bins on its own
            // stream, the transforms numbered bin-major, so that the first inverse chunks (which then need only the first bins'
            // spectra) start when the first group is done and the later groups run beside them -- the step is bound by the sum of
            // the kernels' work, not by the forward side's latency, so this buys nothing measurable and stays off.
            const int H = (int)g.members.size();
            int fwd_groups = std::max(1, std::min(std::min(8, bins), s->tune.fwd_groups));
            if (s->kt.enabled || fused_cfg || lanes_cfg > 2 || T <= chunkT) fwd_groups = 1;
            const bool bin_major = fwd_groups > 1;
            const int gbins = (bins + fwd_groups - 1) / fwd_groups;
            if (bin_major) fwd_groups = (bins + gbins - 1) / gbins;
            if (!bin_major)
                GB_CUDA(s, gb_acq_forward(plan, fp, bins, true, s->stream, &l, s));
            else
                {
                    cudaStream_t fs = s->stream34[0];
                    GB_CUDA(s, cudaEventRecord(s->ev_fork, s->stream));
                    GB_CUDA(s, cudaStreamWaitEvent(fs, s->ev_fork, 0));
                    for (int gi = 0; gi < fwd_groups; gi++)
                        {
                            fp.first_bin = gi * gbins;
                            GB_CUDA(s, gb_acq_forward(plan, fp, std::min(gbins, bins - fp.first_bin), true, fs, &l, s));
                            GB_CUDA(s, cudaEventRecord(s->ev_fwd[gi], fs));
                        }
                }
            bool fused_done = false;
            if (fused_cfg && !g.grid)
                {
                    const int ring = std::max(2, std::min(fused_ring, T));
                    const int lag = std::min(fused_lag, ring - 1);
                    void *d_ring, *d_ctl;
                    if ((rc = s->scratch(17, sizeof(float2) * (size_t)N * ring, &d_ring))) return rc;
                    if ((rc = s->scratch(16, std::max<size_t>(gb_acq_fused_ctl_bytes(T), 16), &d_ctl))) return rc;
                    GbAcqInvParams ip;
                    ip.spectra = (const float2*)d_spec;
                    ip.code_fft = (const float2* const*)((char*)d_up + g.off_code);
                    ip.rowbuf = (float2*)d_ring;
                    ip.n_bins = bins;
                    ip.first_transform = 0;
                    ip.n_transforms = T;
                    ip.keep_offset = a0->c.bit_transition_flag ? N / 2 : 0;
                    ip.keep_len = (int)a0->eff;
                    ip.grid = nullptr;
                    ip.accumulate = 0;
                    ip.partial = (GbPeak*)d_part;
                    ip.tmpbuf = nullptr;
                    ip.scale = nullptr;
                    cudaError_t fe = gb_acq_inverse_fused(plan, ip, d_ctl, lag, ring, s->sm_count, s->stream, &l, s);
                    if (fe == cudaSuccess)
                        fused_done = true;
                    else if (fe != cudaErrorNotSupported)
                        GB_CUDA(s, fe);
                    else
                        (void)cudaGetLastError();
                }
            const bool fork = !fused_done && lanes > 1 && T > chunkT;
            cudaStream_t lane_stream[4] = {s->stream, s->stream2, s->stream34[0], s->stream34[1]};
            cudaEvent_t lane_join[4] = {nullptr, s->ev_join, s->ev_join34[0], s->ev_join34[1]};
            if (fork)
                {
                    GB_CUDA(s, cudaEventRecord(s->ev_fork, s->stream));
                    for (int k = 1; k < lanes; k++) GB_CUDA(s, cudaStreamWaitEvent(lane_stream[k], s->ev_fork, 0));
                }
            int lane = 0;
            int lane_seen[4] = {-1, -1, -1, -1};
            for (int t0 = fused_done ? T : 0; t0 < T; t0 += chunkT, lane = (lane + 1) % lanes)
                {
                    const int ln = fork ? lane : 0;
                    const int nT = std::min(chunkT, T - t0);
                    if (bin_major)
                        {
                            const int need = std::min(fwd_groups - 1, ((t0 + nT - 1) / H) / gbins);  // last forward group this chunk reads
                            if (need > lane_seen[ln])
                                {
                                    GB_CUDA(s, cudaStreamWaitEvent(lane_stream[ln], s->ev_fwd[need], 0));
                                    lane_seen[ln] = need;
                                }
                        }
                    GbAcqInvParams ip;
                    ip.spectra = (const float2*)d_spec;
                    ip.code_fft = (const float2* const*)((char*)d_up + g.off_code);
                    ip.rowbuf = (float2*)d_rows[ln];
                    ip.n_bins = bins;
                    ip.first_transform = t0;
                    ip.n_transforms = nT;
                    ip.n_handles = bin_major ? H : 0;
                    ip.keep_offset = a0->c.bit_transition_flag ? N / 2 : 0;
                    ip.keep_len = (int)a0->eff;
                    ip.grid = g.grid ? (float* const*)((char*)d_up + g.off_grid) : nullptr;
                    ip.accumulate = (g.grid && a0->dwell_counter >= 1) ? 1 : 0;  // this is dwell number dwell_counter+1
                    ip.partial = (GbPeak*)d_part;
                    ip.tmpbuf = (g.grid && !a0->use_cfar && !a0->scaled_accum) ? (float* const*)((char*)d_up + g.off_tmp) : nullptr;
                    ip.scale = (g.grid && a0->scaled_accum == 1) ? d_power + 1 : nullptr;
                    GB_CUDA(s, gb_acq_inverse(plan, ip, lane_stream[ln], &l, s));
                }
            if (fork)
                for (int k = 1; k < lanes; k++)
                    {
                        GB_CUDA(s, cudaEventRecord(lane_join[k], lane_stream[k]));
                        GB_CUDA(s, cudaStreamWaitEvent(s->stream, lane_join[k], 0));
                    }
            else if (power_aside)
                {
                    GB_CUDA(s, cudaEventRecord(s->ev_join, s->stream2));
                    GB_CUDA(s, cudaStreamWaitEvent(s->stream, s->ev_join, 0));
                }
            GbPeak* d_peaks = (GbPeak*)((char*)d_res + g.off_peaks);
            // every mode leaves per-tile maxima in d_part (over the stored grid where one is kept)
            GB_CUDA(s, gb_acq_reduce_peaks((const GbPeak*)d_part, (int)g.members.size(), bins, tiles, d_peaks, s->stream, &l));
            // first/second-peak statistic (:633-660): the second peak of every member's winning row, one CTA each, in this
            // launch phase -- the results ride back with the peaks
            if (g.grid && !a0->use_cfar && !a0->scaled_accum)
                GB_CUDA(s, gb_acq_second_peak_batch((float* const*)((char*)d_up + g.off_grid), (float* const*)((char*)d_up + g.off_tmp), d_peaks,
                               (int)g.members.size(), (int)a0->eff, N, (int)a0->c.samples_per_chip, (float*)((char*)d_res + g.off_second), s->stream, &l));
        }
    s->launches += l;
    { int rc0 = rings_release(s); if (rc0) return rc0; }
    GB_CUDA(s, cudaMemcpyAsync(ctx.h_res, d_res, ctx.res_bytes, cudaMemcpyDeviceToHost, s->stream));
    for (int i = 0; i < n; i++)
        {
            list[i]->pending = true;
            list[i]->pending_first = first_sample_index;
        }
    for (size_t gi = 0; gi < ctx.groups.size(); gi++)
        for (size_t k = 0; k < ctx.groups[gi].members.size(); k++)
            {
                list[ctx.groups[gi].members[k]]->pending_group = (int)gi;
                list[ctx.groups[gi].members[k]]->pending_slot = (int)k;
            }
    ctx.active = true;
    ctx.sig.swap(sig);
    return GB200_OK;
}

// decision logic of acquisition_core, pcps_acquisition.cc:832-929 (events -> `decision`)
static void acq_decide(gb200_acq* a, float stat, gb200_acq_result* out)
{
    bool positive = false, negative = false;
    int state = 1;
    if (!a->c.bit_transition_flag)
        {
            if (stat > a->c.threshold)
                {
                    if (a->c.make_2_steps)
                        {
                            if (a->step_two)
                                {
                                    positive = true;
                                    a->step_two = false;
                                    state = 0;
                                }
                            else
                                {
                                    a->step_two = true;
                                    a->dwell_counter = 0;
                                    a->positive_acq = 0;
                                    state = 0;
                                }
                        }
                    else
                        {
                            positive = true;
                            state = 0;
                        }
                }
            else
                state = 1;
            if (a->dwell_counter == a->c.max_dwells)
                {
                    if (state != 0) negative = true;
                    state = 0;
                    a->step_two = false;
                }
        }
    else
        {
            if (stat > a->c.threshold)
                {
                    if (a->c.make_2_steps)
                        {
                            if (a->step_two)
                                {
                                    positive = true;
                                    a->step_two = false;
                                }
                            else
                                {
                                    a->step_two = true;
                                    a->dwell_counter = 0;
                                }
                        }
                    else
                        positive = true;
                }
            else
                {
                    a->step_two = false;
                    negative = true;
                }
        }
    if (positive) a->positive_acq = 1;
    if (negative) a->positive_acq = 0;
    out->decision = positive ? 1 : (negative ? 0 : (a->step_two ? 3 : 2));
    if (a->step_two) a->doppler_center_step_two = (float)out->acq_doppler_hz;  // general_work :964-970
}

extern "C" int gb200_acq_collect_batch(gb200_acq* const* list, int n, gb200_acq_result* out)
{
    if (!list || n <= 0 || !list[0] || !out) return GB200_E_INVALID;
    gb200_service* s = list[0]->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    AcqBatchCtx& ctx = s->acq_ctx;
    if (!ctx.active)
        {
            s->set_error("gb200_acq_collect_batch: nothing launched");
            return GB200_E_STATE;
        }
    GB_CUDA(s, cudaStreamSynchronize(s->stream));
    const char* h_res = ctx.h_res;
    int rc = GB200_OK;
    for (int i = 0; i < n; i++)
        {
            gb200_acq* a = list[i];
            if (!a->pending || a->pending_group < 0 || a->pending_group >= (int)ctx.groups.size())
                {
                    s->set_error("gb200_acq_collect_batch: handle was not part of the launched batch");
                    ctx.active = false;
                    return GB200_E_STATE;
                }
            const AcqGroup& g = ctx.groups[a->pending_group];
            const float power = *(const float*)(h_res + g.off_power);
            const GbPeak pk = ((const GbPeak*)(h_res + g.off_peaks))[a->pending_slot];
            gb200_acq_result r;
            memset(&r, 0, sizeof r);
            a->dwell_counter++;
            const bool was_step_two = a->step_two;
            r.code_phase_index = pk.idx;
            r.doppler_bin = pk.bin;
            r.peak_mag = pk.mag;
            if (!was_step_two)
                r.doppler_hz = -(int32_t)a->c.doppler_max + (int32_t)a->c.doppler_step * (int32_t)pk.bin;  // :585-588
            else
                r.doppler_hz = (int32_t)(a->doppler_center_step_two + ((float)pk.bin - (float)floor(a->c.num_doppler_bins_step2 / 2.0)) * a->c.doppler_step2);  // :591
            float stat;
            if (a->scaled_accum)
                {
                    // Tong detector: the kept grid already holds the accumulated test statistic; the caller owns the
                    // counter logic and ends the attempt with gb200_acq_reset
                    r.input_power = power;
                    r.test_statistic = pk.mag;
                    r.acq_delay_samples = (double)fmodf((float)pk.idx, a->c.samples_per_code);
                    r.acq_doppler_hz = (double)r.doppler_hz;
                    r.acq_samplestamp = a->pending_first + a->consumed;
                    r.decision = 2;
                    a->pending = false;
                    out[i] = r;
                    continue;
                }
            if (a->use_cfar)
                {
                    r.input_power = power;
                    const float nf = (float)a->fft_size * (float)a->fft_size;  // :571
                    const float magt = pk.mag / (nf * nf);
                    stat = magt / power;
                }
            else
                {
                    r.input_power = a->c.bit_transition_flag ? power : 0.0f;
                    // second peak of the winning row (:633-660), found on the device in the launch phase
                    const float second = ((const float*)(h_res + g.off_second))[a->pending_slot];
                    stat = pk.mag / second;
                }
            r.test_statistic = stat;
            r.acq_delay_samples = (double)fmodf((float)pk.idx, a->c.samples_per_code);  // :766 (float32 fmod)
            r.acq_doppler_hz = (double)r.doppler_hz;
            r.acq_samplestamp = a->pending_first + a->consumed;  // d_sample_counter after buffering, SURVEY 3.2
            acq_decide(a, stat, &r);
            // end-of-attempt reset, :912-929
            if (a->dwell_counter == a->c.max_dwells || a->positive_acq == 1)
                {
                    a->dwell_counter = 0;
                    a->positive_acq = 0;
                    if (a->d_grid && !a->keep_grid) a->grid_zero_pending = true;
                }
            a->pending = false;
            out[i] = r;
        }
    ctx.active = false;
    return rc;
}

// = AcquisitionInterface::stop_acquisition() for a search that is in flight: the launched batch is abandoned -- nothing is
// waited for, no result is read, no handle's dwell / step-two state advances (the kernels still run to completion on the stream)
extern "C" int gb200_acq_discard_batch(gb200_acq* const* list, int n)
{
    if (!list || n <= 0 || !list[0]) return GB200_E_INVALID;
    gb200_service* s = list[0]->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    AcqBatchCtx& ctx = s->acq_ctx;
    if (!ctx.active)
        {
            s->set_error("gb200_acq_discard_batch: nothing launched");
            return GB200_E_STATE;
        }
    for (int i = 0; i < n; i++)
        if (list[i]) list[i]->pending = false;
    ctx.active = false;
    return GB200_OK;
}

extern "C" int gb200_acq_run_batch(gb200_acq* const* list, int n, int stream_id, uint64_t first_sample_index, gb200_acq_result* out)
{
    if (!list || n <= 0 || !list[0]) return GB200_E_INVALID;
    // ONE critical section for launch + collect: another channel's thread (or a set_local_code) cannot get between the two
    std::lock_guard<std::recursive_mutex> lk(list[0]->s->mu);
    int rc = gb200_acq_launch_batch(list, n, stream_id, first_sample_index);
    if (rc) return rc;
    return gb200_acq_collect_batch(list, n, out);
}

extern "C" int gb200_acq_run(gb200_acq* a, int stream_id, uint64_t first_sample_index, gb200_acq_result* out)
{
    if (!a || !out) return GB200_E_INVALID;
    gb200_service* s = a->s;
    if (s->batch_window_us == 0)
        {
            gb200_acq* l[1] = {a};
            return gb200_acq_run_batch(l, 1, stream_id, first_sample_index, out);
        }
    // ---- coalescing: every channel's acquisition thread queues its search; the first one in collects for batch_window_us,
    // then runs ONE batch per (stream, first sample) group and hands each caller its own result ----
    gb200_service::AcqPending me{a, stream_id, first_sample_index, out, GB200_OK, false};
    std::unique_lock<std::mutex> lk(s->aqmu);
    s->acq_queue.push_back(&me);
    if (s->acq_collecting)
        {
            s->aqcv.wait(lk, [&] { return me.done; });
            return me.rc;
        }
    s->acq_collecting = true;
    s->aqcv.wait_for(lk, std::chrono::microseconds(s->batch_window_us), [] { return false; });
    std::vector<gb200_service::AcqPending*> batch;
    batch.swap(s->acq_queue);
    s->acq_collecting = false;  // the next arrival collects the next batch while this one runs (the service mutex serialises the GPU work)
    lk.unlock();
    std::vector<bool> taken(batch.size(), false);
    uint64_t groups = 0;
    for (size_t i = 0; i < batch.size(); i++)
        {
            if (taken[i]) continue;
            std::vector<size_t> members;
            for (size_t k = i; k < batch.size(); k++)
                if (!taken[k] && batch[k]->stream_id == batch[i]->stream_id && batch[k]->first == batch[i]->first)
                    {
                        taken[k] = true;
                        members.push_back(k);
                    }
            std::vector<gb200_acq*> list(members.size());
            std::vector<gb200_acq_result> res(members.size());
            for (size_t k = 0; k < members.size(); k++) list[k] = batch[members[k]]->a;
            const int rc = gb200_acq_run_batch(list.data(), (int)list.size(), batch[i]->stream_id, batch[i]->first, res.data());
            if (rc != GB200_OK && members.size() > 1)
                {
                    // one bad request must not fail the others: redo one by one.  The batch failed in its launch phase (collect has
                    // no failing step once the results are back), so no handle's dwell state has advanced yet.
                    for (size_t k = 0; k < members.size(); k++)
                        batch[members[k]]->rc = gb200_acq_run_batch(&list[k], 1, batch[i]->stream_id, batch[i]->first, batch[members[k]]->out);
                }
            else
                for (size_t k = 0; k < members.size(); k++)
                    {
                        batch[members[k]]->rc = rc;
                        if (rc == GB200_OK) *batch[members[k]]->out = res[k];
                    }
            groups++;
        }
    lk.lock();
    s->acq_coalesced_batches += groups;
    s->acq_coalesced_requests += batch.size();
    for (auto* b : batch) b->done = true;
    const int my_rc = me.rc;
    lk.unlock();
    s->aqcv.notify_all();
    return my_rc;
}

extern "C" int gb200_service_acq_coalescing_stats(gb200_service* s, uint64_t* batches, uint64_t* requests)
{
    if (!s) return GB200_E_INVALID;
    std::lock_guard<std::mutex> lk(s->aqmu);
    if (batches) *batches = s->acq_coalesced_batches;
    if (requests) *requests = s->acq_coalesced_requests;
    return GB200_OK;
}

// ------------------------------------------------------------------ tracking
struct gb200_trk
{
    gb200_service* s = nullptr;
    gb200_trk_cfg c;
    float* d_code = nullptr;       // float[code_len] or float2[code_len]
    float* d_data_code = nullptr;  // float[code_len] or nullptr
    // "extended table": the code repeated over ext_entries window entries in the layout the kernel stages (float, or
    // float2 = (pilot, data) / (re, im)), so that a CTA's whole code window is one contiguous 16-byte-aligned range
    float* d_ext = nullptr;
    int ext_entries = 0;
    std::vector<float> h_code, h_data;  // host copies the extended table is rebuilt from
    int code_len = 0;
    int n_taps = 0;
    int prompt_tap = 0;
    bool has_data = false;
    float shifts[GB200_MAX_TAPS];
    // resources of the single-request path (gb200_trk_correlate)
    gb200_trk_batch* single = nullptr;
};

struct TrkGroup
{
    int first = 0, count = 0;  // range in the (sorted) work array
    int n_taps = 0, item_type = 0, code_kind = 0, max_n = 0, prompt_tap = 0;
    bool has_data = false, high_dyn = false;
    cudaEvent_t ready = nullptr;  // push event the group's requests wait for (the push that delivered their newest sample)
    uint64_t ready_seq = 0;
};

struct gb200_trk_batch
{
    gb200_service* s = nullptr;
    int n = 0;
    // Request records are staged in pinned, device-mapped host memory and pulled to the device by a small kernel: a
    // cudaMemcpyAsync would queue on the host-to-device copy engine BEHIND the sample pushes still in flight and hold the
    // first block's launch back until the last block has landed.
    struct PinnedWork
    {
        GbTrkWork* p = nullptr;
        GbTrkWork* dev = nullptr;  // device alias of p
        size_t cap = 0, n = 0;
        GbTrkWork& operator[](size_t i) { return p[i]; }
        GbTrkWork* data() { return p; }
        bool resize(size_t m)
        {
            if (m > cap)
                {
                    if (p) cudaFreeHost(p);
                    p = nullptr;
                    cap = 0;
                    const size_t want = std::max<size_t>(m, 64);
                    if (cudaHostAlloc((void**)&p, sizeof(GbTrkWork) * want, cudaHostAllocMapped) != cudaSuccess) return false;
                    if (cudaHostGetDevicePointer((void**)&dev, p, 0) != cudaSuccess) return false;
                    cap = want;
                }
            n = m;
            return true;
        }
        ~PinnedWork()
        {
            if (p) cudaFreeHost(p);
        }
    } h_work;
    std::vector<int> order;  // work slot -> request index
    std::vector<TrkGroup> groups;
    GbTrkWork* d_work = nullptr;
    float2* d_out = nullptr;
    float2* d_partials = nullptr;
    unsigned* d_counters = nullptr;
    float2* h_out = nullptr;  // pinned
    int chunk = 4096;
    int win_entries = 1032;  // code-window entries per CTA (sizes the dynamic shared memory)
    int max_chunks = 1;
};

extern "C" int gb200_trk_create(gb200_service* s, const gb200_trk_cfg* cfg, gb200_trk** out)
{
    if (!s || !cfg || !out) return GB200_E_INVALID;
    *out = nullptr;
    if (cfg->n_taps < 1 || cfg->n_taps > GB200_MAX_TAPS || cfg->max_code_len < 1 || cfg->item_type < 0 || cfg->item_type > 2 ||
        (cfg->code_kind != GB200_CODE_REAL && cfg->code_kind != GB200_CODE_CPLX && cfg->code_kind != GB200_CODE_16SC) ||
        (cfg->high_dyn && cfg->code_kind != GB200_CODE_REAL) || (cfg->code_kind == GB200_CODE_16SC && cfg->item_type != GB200_ITEM_16IC))
        {
            s->set_error("gb200_trk_create: bad configuration");
            return GB200_E_INVALID;
        }
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    gb200_trk* t = new gb200_trk();
    t->s = s;
    t->c = *cfg;
    const size_t w = cfg->code_kind == GB200_CODE_CPLX ? 2 : 1;
    cudaError_t e = cudaMalloc(&t->d_code, sizeof(float) * w * cfg->max_code_len);
    if (e == cudaSuccess && cfg->code_kind == GB200_CODE_REAL) e = cudaMalloc(&t->d_data_code, sizeof(float) * cfg->max_code_len);
    if (e == cudaSuccess && cfg->code_kind != GB200_CODE_16SC) e = cudaMalloc(&t->d_ext, sizeof(float) * 2 * (2 * (size_t)cfg->max_code_len + 2048));
    if (e != cudaSuccess)
        {
            s->set_error(std::string("cudaMalloc code: ") + cudaGetErrorString(e));
            if (t->d_code) cudaFree(t->d_code);
            delete t;
            return GB200_E_CUDA;
        }
    if (cfg->code_kind != GB200_CODE_16SC)
        {
            std::lock_guard<std::mutex> ql(s->qmu);
            s->live_trk++;
        }
    *out = t;
    return GB200_OK;
}

extern "C" int gb200_trk_destroy(gb200_trk* t)
{
    if (!t) return GB200_E_INVALID;
    gb200_service* s = t->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    cudaStreamSynchronize(s->stream);
    if (t->c.code_kind != GB200_CODE_16SC)
        {
            std::lock_guard<std::mutex> ql(s->qmu);
            s->live_trk--;
        }
    if (t->single) gb200_trk_batch_destroy(t->single);
    if (t->d_code) cudaFree(t->d_code);
    if (t->d_data_code) cudaFree(t->d_data_code);
    if (t->d_ext) cudaFree(t->d_ext);
    delete t;
    return GB200_OK;
}

// (re)build the extended table from the host copies of the code(s); caller holds s->mu
static int trk_rebuild_ext(gb200_trk* t)
{
    gb200_service* s = t->s;
    if (!t->d_ext || t->code_len < 1) return GB200_OK;
    const int L = t->code_len;
    const bool cplx = t->c.code_kind == GB200_CODE_CPLX;
    const int ew = (cplx || t->has_data) ? 2 : 1;
    const int entries = (2 * L + 2048) & ~3;
    std::vector<float> ext((size_t)entries * ew);
    for (int j = 0; j < entries; j++)
        {
            const int k = j % L;
            if (cplx)
                {
                    ext[2 * (size_t)j] = t->h_code[2 * (size_t)k];
                    ext[2 * (size_t)j + 1] = t->h_code[2 * (size_t)k + 1];
                }
            else if (t->has_data)
                {
                    ext[2 * (size_t)j] = t->h_code[k];
                    ext[2 * (size_t)j + 1] = t->h_data[k];
                }
            else
                ext[j] = t->h_code[k];
        }
    t->ext_entries = 0;
    GB_CUDA(s, cudaStreamSynchronize(s->stream));  // launches in flight may still read the old table
    GB_CUDA(s, cudaMemcpy(t->d_ext, ext.data(), sizeof(float) * ext.size(), cudaMemcpyHostToDevice));
    t->ext_entries = entries;
    return GB200_OK;
}

extern "C" int gb200_trk_set_local_code_and_taps(gb200_trk* t, const void* code, int code_len, const float* shifts, int n_taps)
{
    if (!t || !code || !shifts) return GB200_E_INVALID;
    gb200_service* s = t->s;
    if (code_len < 1 || code_len > t->c.max_code_len || n_taps < 1 || n_taps > t->c.n_taps)
        {
            s->set_error("gb200_trk_set_local_code_and_taps: code_len / n_taps exceed what the handle was created for");
            return GB200_E_INVALID;
        }
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    const size_t w = t->c.code_kind == GB200_CODE_CPLX ? 2 : 1;
    GB_CUDA(s, cudaMemcpyAsync(t->d_code, code, sizeof(float) * w * code_len, cudaMemcpyHostToDevice, s->stream));
    GB_CUDA(s, cudaStreamSynchronize(s->stream));  // `code` is caller memory: do not keep reading it
    if (t->has_data && (code_len != t->code_len || t->prompt_tap >= n_taps)) t->has_data = false;  // a data table of another length, or one whose prompt tap no longer exists, cannot ride along
    t->code_len = code_len;
    t->n_taps = n_taps;
    for (int i = 0; i < GB200_MAX_TAPS; i++) t->shifts[i] = i < n_taps ? shifts[i] : 0.f;
    if (t->c.code_kind != GB200_CODE_16SC)
        {
            t->h_code.assign((const float*)code, (const float*)code + w * code_len);
            return trk_rebuild_ext(t);
        }
    return GB200_OK;
}

extern "C" int gb200_trk_set_taps(gb200_trk* t, const float* shifts, int n_taps)
{
    if (!t || !shifts) return GB200_E_INVALID;
    gb200_service* s = t->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    if (n_taps != t->n_taps)
        {
            s->set_error("gb200_trk_set_taps: n_taps differs from the handle's current tap count");
            return GB200_E_INVALID;
        }
    for (int i = 0; i < n_taps; i++) t->shifts[i] = shifts[i];
    return GB200_OK;
}

extern "C" int gb200_trk_set_data_code(gb200_trk* t, const float* code, int code_len, int prompt_tap)
{
    if (!t) return GB200_E_INVALID;
    gb200_service* s = t->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    if (!code)
        {
            t->has_data = false;
            cudaSetDevice(s->device);
            return trk_rebuild_ext(t);
        }
    if (t->c.code_kind != GB200_CODE_REAL || code_len != t->code_len || prompt_tap < 0 || prompt_tap >= t->n_taps)
        {
            s->set_error("gb200_trk_set_data_code: needs a real-code handle, the tracking code's length and a valid prompt tap");
            return GB200_E_INVALID;
        }
    cudaSetDevice(s->device);
    GB_CUDA(s, cudaMemcpyAsync(t->d_data_code, code, sizeof(float) * code_len, cudaMemcpyHostToDevice, s->stream));
    GB_CUDA(s, cudaStreamSynchronize(s->stream));
    t->has_data = true;
    t->prompt_tap = prompt_tap;
    t->h_data.assign(code, code + code_len);
    return trk_rebuild_ext(t);
}

static int trk_fill_work(gb200_service* s, const gb200_trk_req& r, int out_index, GbTrkWork* w)
{
    gb200_trk* t = r.trk;
    if (!t || t->s != s || t->code_len == 0 || r.n_samples < 1)
        {
            s->set_error("tracking request: handle not initialised (set_local_code_and_taps) or n_samples < 1");
            return GB200_E_STATE;
        }
    if (t->c.code_kind == GB200_CODE_16SC)
        {
            s->set_error("tracking request: GB200_CODE_16SC handles are served by gb200_trk_correlate_16sc[_batch]");
            return GB200_E_INVALID;
        }
    if ((uint64_t)r.n_samples >= (1u << 24))
        {
            s->set_error("tracking request: n_samples must be < 2^24");
            return GB200_E_INVALID;
        }
    const void* win;
    int rc = ring_window(s, r.stream_id, r.first_sample_index, (uint64_t)r.n_samples, t->c.item_type, &win);
    if (rc) return rc;
    memset(w, 0, sizeof(*w));
    w->samples = win;
    w->code = t->d_code;
    w->data_code = t->has_data ? t->d_data_code : nullptr;
    w->ext = t->ext_entries > 0 ? t->d_ext : nullptr;
    w->ext_entries = t->ext_entries;
    w->code_len = t->code_len;
    w->n = r.n_samples;
    w->n_taps = t->n_taps;
    w->prompt_tap = t->prompt_tap;
    memcpy(w->shifts, t->shifts, sizeof(w->shifts));
    w->rem_carr = r.rem_carr_phase_rad;
    w->phase_step = r.phase_step_rad;
    w->phase_rate = r.phase_rate_step_rad;
    w->rem_code = r.rem_code_phase_chips;
    w->code_step = r.code_phase_step_chips;
    w->code_rate = r.code_phase_rate_step_chips;
    w->out_index = out_index;
    {
        // the reference's float32 NCO constants (cpu_multicorrelator_real_codes.cc:140-149), host libm like the reference
        const float p0r = cosf(r.rem_carr_phase_rad), p0i = -sinf(r.rem_carr_phase_rad);
        const float incr = cosf(-r.phase_step_rad), inci = sinf(-r.phase_step_rad);
        const double inv2pi = 0.15915494309189533576888376337251;
        w->nco_a0_turns = atan2((double)p0i, (double)p0r) * inv2pi;
        w->nco_theta_turns = atan2((double)inci, (double)incr) * inv2pi;
        const double mag = hypot((double)incr, (double)inci);
        w->nco_mag_delta = (float)(mag - 1.0);
        w->nco_log_mag = log(mag);
        {
            // 64-bit fixed-point turn counts for the kernels (wrap mod 1 turn in integer arithmetic)
            auto to_fix = [](double turns) -> unsigned long long {
                double f = turns - floor(turns);  // [0, 1)
                if (f >= 1.0) f = 0.0;
                const double hi = floor(f * 4294967296.0);
                const double lo = floor((f * 4294967296.0 - hi) * 4294967296.0);
                return ((unsigned long long)hi << 32) | (unsigned long long)std::min(lo, 4294967295.0);
            };
            w->nco_a0_fix = to_fix(w->nco_a0_turns);
            w->nco_theta_fix = to_fix(w->nco_theta_turns);
        }
    }
    return GB200_OK;
}

static void trk_group_key(const gb200_trk* t, TrkGroup* g)
{
    g->n_taps = t->n_taps;
    g->item_type = t->c.item_type;
    g->code_kind = t->c.code_kind;
    g->has_data = t->has_data;
    g->high_dyn = t->c.high_dyn != 0;
    g->prompt_tap = t->prompt_tap;
}

static bool trk_same_group(const TrkGroup& a, const TrkGroup& b)
{
    return a.n_taps == b.n_taps && a.item_type == b.item_type && a.code_kind == b.code_kind && a.has_data == b.has_data && a.high_dyn == b.high_dyn &&
           (!a.has_data || a.prompt_tap == b.prompt_tap) && a.ready == b.ready;
}

// (re)build the work array of a batch from requests; buffers must already be sized for b->n requests
// host threads that fill a long batch of request records (tuning key "host_threads", default 8)
// -- never more than the cores this process may run on (a rank pinned to 4 cores with 8 staging threads spends its time in the
// OpenMP barrier: 14 ms instead of 0.9 ms per 4160-request batch at 8 ranks per host)
[[maybe_unused]] static int gb_host_threads(const gb200_service* s)
{
    int avail = 64;
    cpu_set_t set;
    if (sched_getaffinity(0, sizeof(set), &set) == 0) avail = std::max(1, CPU_COUNT(&set));
    return std::max(1, std::min(std::min(64, s->tune.host_threads), avail));
}

static int trk_batch_stage(gb200_trk_batch* b, const gb200_trk_req* reqs)
{
    gb200_service* s = b->s;
    const int n = b->n;
    // bucket requests by kernel variant, keeping request order inside a bucket
    b->groups.clear();
    const bool trace = s->tune.trace != 0;
    const auto ts0 = std::chrono::steady_clock::now();
    std::vector<int> gid(n);
    std::vector<TrkGroup> keys(n);
    for (int i = 0; i < n; i++)
        if (!reqs[i].trk) return GB200_E_INVALID;
#pragma omp parallel for schedule(static) num_threads(gb_host_threads(s)) if (n >= 1024)
    for (int i = 0; i < n; i++)
        {
            TrkGroup& k = keys[i];
            trk_group_key(reqs[i].trk, &k);
            if (reqs[i].stream_id >= 0 && reqs[i].stream_id < (int)s->rings.size())
                if (const GbRing::Mark* m = s->rings[reqs[i].stream_id].mark_for(reqs[i].first_sample_index + (uint64_t)std::max(reqs[i].n_samples, 0)))
                    {
                        k.ready = m->ev;
                        k.ready_seq = m->seq;
                    }
        }
    for (int i = 0; i < n; i++)
        {
            const TrkGroup& k = keys[i];
            int found = -1;
            if (i > 0 && trk_same_group(b->groups[gid[i - 1]], k))
                found = gid[i - 1];  // the common case: consecutive requests of one channel
            else
                for (size_t g = 0; g < b->groups.size(); g++)
                    if (trk_same_group(b->groups[g], k)) found = (int)g;
            if (found < 0)
                {
                    b->groups.push_back(k);
                    found = (int)b->groups.size() - 1;
                }
            gid[i] = found;
            b->groups[found].count++;
            b->groups[found].max_n = std::max(b->groups[found].max_n, reqs[i].n_samples);
        }
    {
        // launch order = push order, so an early block's kernel never queues behind a wait for a later block
        std::vector<int> perm(b->groups.size()), inv(b->groups.size());
        for (size_t g = 0; g < perm.size(); g++) perm[g] = (int)g;
        std::stable_sort(perm.begin(), perm.end(), [&](int x, int y) { return b->groups[x].ready_seq < b->groups[y].ready_seq; });
        std::vector<TrkGroup> sorted;
        for (size_t g = 0; g < perm.size(); g++)
            {
                sorted.push_back(b->groups[perm[g]]);
                inv[perm[g]] = (int)g;
            }
        b->groups.swap(sorted);
        for (int i = 0; i < n; i++) gid[i] = inv[gid[i]];
    }
    int off = 0;
    for (auto& g : b->groups)
        {
            g.first = off;
            off += g.count;
            g.count = 0;
        }
    if (!b->h_work.resize(n))
        {
            s->set_error("tracking batch: pinned staging allocation failed");
            return GB200_E_CUDA;
        }
    b->order.resize(n);
    const auto ts1 = std::chrono::steady_clock::now();
    int max_n = 1;
    std::vector<int> slot_of(n);
    for (int i = 0; i < n; i++)
        {
            TrkGroup& g = b->groups[gid[i]];
            slot_of[i] = g.first + g.count++;
            b->order[slot_of[i]] = i;
            max_n = std::max(max_n, reqs[i].n_samples);
        }
    {
        // the per-request NCO constants go through host libm in float32 and float64 (~0.1 us each): a long batch is filled
        // by a few host threads
        int first_rc = GB200_OK;
#pragma omp parallel for schedule(static) num_threads(gb_host_threads(s)) if (n >= 1024)
        for (int i = 0; i < n; i++)
            {
                const int rc = trk_fill_work(s, reqs[i], i, &b->h_work[slot_of[i]]);
                if (rc)
                    {
#pragma omp critical(gb200_fill_rc)
                        if (!first_rc) first_rc = rc;
                    }
            }
        if (first_rc) return first_rc;
    }
    if (trace)
        {
            const auto ts2 = std::chrono::steady_clock::now();
            auto us = [](auto a, auto z) { return (double)std::chrono::duration_cast<std::chrono::nanoseconds>(z - a).count() * 1e-3; };
            fprintf(stderr, "  stage: keys+groups %.1f us, fill %.1f us\n", us(ts0, ts1), us(ts1, ts2));
        }
    // chunk size (samples per CTA): the largest that still gives >= 4 CTAs per SM and whose code window (the chips one
    // CTA can touch: chunk*step + tap spread) fits the shared-memory staging area; per-CTA set-up (NCO seed, window
    // fill, 6-accumulator reduction) is ~700 instructions per thread, so long chunks matter.
    double chips_per_sample = 0.0, spread = 0.0;
    for (int i = 0; i < n; i++)
        {
            const GbTrkWork& w = b->h_work[i];
            float lo = w.shifts[0], hi = w.shifts[0];
            for (int t = 1; t < w.n_taps; t++)
                {
                    lo = std::min(lo, w.shifts[t]);
                    hi = std::max(hi, w.shifts[t]);
                }
            spread = std::max(spread, (double)hi - (double)lo);
            chips_per_sample = std::max(chips_per_sample, std::fabs((double)w.code_step));
        }
    // window entries are 8 bytes when they carry (pilot, data) chips or complex chips, else 4
    bool wide = false;
    for (const auto& g : b->groups) wide = wide || g.has_data || g.code_kind == GB200_CODE_CPLX;
    // + 1023 samples: the kernel pads every thread's work to whole groups of four (4 x 256 samples)
    auto win_for = [&](int c) { return ((int)ceil((double)(std::min(c, max_n) + 1023) * chips_per_sample + spread) + 8 + 1) & ~1; };
    long long ctas = 0;
    b->chunk = 1024;
    const int chunk_cap = s->tune.trk_chunk_max;  // tuning only
    for (int c : {32768, 16384, 8192, 4096, 2048, 1024})
        {
            if (c > chunk_cap && c > 1024) continue;
            b->chunk = c;
            // prefer windows that leave room for 3 CTAs per SM; the smallest chunk is always accepted (a window that does
            // not fit even then makes the kernel take its table-in-L2 path)
            if (c > 1024 && gb_trk_smem_bytes(win_for(c), wide, c) > (size_t)GB200_TRK_SMEM_PREF) continue;
            ctas = 0;
            for (int i = 0; i < n; i++) ctas += (reqs[i].n_samples + c - 1) / c;
            if (ctas >= 4LL * s->sm_count) break;
        }
    b->win_entries = win_for(b->chunk);
    while (b->win_entries > 2 && gb_trk_smem_bytes(b->win_entries, wide, b->chunk) > (size_t)GB200_TRK_SMEM_MAX) b->win_entries -= 2;
    b->max_chunks = (max_n + b->chunk - 1) / b->chunk;
    return GB200_OK;
}

static int trk_batch_alloc(gb200_service* s, int n, int max_chunks_cap, gb200_trk_batch** out)
{
    gb200_trk_batch* b = new gb200_trk_batch();
    b->s = s;
    b->n = n;
    cudaError_t e = cudaMalloc(&b->d_work, sizeof(GbTrkWork) * n);
    if (e == cudaSuccess) e = cudaMalloc(&b->d_out, sizeof(float2) * GB200_OUT_STRIDE * n);
    if (e == cudaSuccess) e = cudaMalloc(&b->d_partials, sizeof(float2) * GB200_OUT_STRIDE * (size_t)n * max_chunks_cap);
    if (e == cudaSuccess) e = cudaMalloc(&b->d_counters, sizeof(unsigned) * n);
    if (e == cudaSuccess) e = cudaMemsetAsync(b->d_counters, 0, sizeof(unsigned) * n, s->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(b->d_out, 0, sizeof(float2) * GB200_OUT_STRIDE * n, s->stream);
    if (e == cudaSuccess) e = cudaMallocHost((void**)&b->h_out, sizeof(float2) * GB200_OUT_STRIDE * n);
    if (e != cudaSuccess)
        {
            s->set_error(std::string("tracking batch allocation: ") + cudaGetErrorString(e));
            gb200_trk_batch_destroy(b);
            return GB200_E_CUDA;
        }
    *out = b;
    return GB200_OK;
}

extern "C" int gb200_trk_batch_create(gb200_service* s, int n, const gb200_trk_req* reqs, gb200_trk_batch** out)
{
    if (!s || n <= 0 || !reqs || !out) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    *out = nullptr;
    int max_n = 1;
    for (int i = 0; i < n; i++) max_n = std::max(max_n, reqs[i].n_samples);
    gb200_trk_batch* b;
    int rc = trk_batch_alloc(s, n, (max_n + 1023) / 1024, &b);
    if (rc) return rc;
    rc = trk_batch_stage(b, reqs);
    if (rc)
        {
            gb200_trk_batch_destroy(b);
            return rc;
        }
    cudaError_t e = gb_copy_words(b->d_work, b->h_work.dev, sizeof(GbTrkWork) * (size_t)n, s->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
    if (e != cudaSuccess)
        {
            s->set_error(std::string("tracking batch upload: ") + cudaGetErrorString(e));
            gb200_trk_batch_destroy(b);
            return GB200_E_CUDA;
        }
    *out = b;
    return GB200_OK;
}

extern "C" int gb200_trk_batch_launch(gb200_trk_batch* b)
{
    if (!b) return GB200_E_INVALID;
    gb200_service* s = b->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    int l = 0;
    for (const auto& g : b->groups)
        {
            if (g.ready) GB_CUDA(s, cudaStreamWaitEvent(s->stream, g.ready, 0));
            // every group shares max_chunks (the partial-sum pitch) so request slots stay independent
            gb_kt_begin(s, GB200_KCLASS_TRK);
            cudaError_t e = gb_trk_launch(b->d_work + g.first, g.count, b->max_chunks * b->chunk, g.n_taps, g.has_data, g.prompt_tap, g.item_type, g.code_kind,
                g.high_dyn, b->d_partials, b->d_counters, b->d_out, b->chunk, b->win_entries, s->stream, &l);
            gb_kt_end(s);
            GB_CUDA(s, e);
        }
    s->launches += l;
    return rings_release(s);
}

extern "C" int gb200_trk_batch_fetch(gb200_trk_batch* b, float* outs_iq)
{
    if (!b || !outs_iq) return GB200_E_INVALID;
    gb200_service* s = b->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    GB_CUDA(s, cudaMemcpyAsync(b->h_out, b->d_out, sizeof(float2) * GB200_OUT_STRIDE * b->n, cudaMemcpyDeviceToHost, s->stream));
    GB_CUDA(s, cudaStreamSynchronize(s->stream));
    memcpy(outs_iq, b->h_out, sizeof(float2) * GB200_OUT_STRIDE * b->n);
    return GB200_OK;
}

extern "C" int gb200_trk_batch_destroy(gb200_trk_batch* b)
{
    if (!b) return GB200_E_INVALID;
    gb200_service* s = b->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    cudaStreamSynchronize(s->stream);
    if (b->d_work) cudaFree(b->d_work);
    if (b->d_out) cudaFree(b->d_out);
    if (b->d_partials) cudaFree(b->d_partials);
    if (b->d_counters) cudaFree(b->d_counters);
    if (b->h_out) cudaFreeHost(b->h_out);
    delete b;
    return GB200_OK;
}

extern "C" int gb200_trk_correlate_batch(gb200_service* s, int n, const gb200_trk_req* reqs, float* outs_iq)
{
    if (!s || n <= 0 || !reqs || !outs_iq) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    // grow-only arena owned by the service: repeated calls (one per block of epochs) do no allocation
    int max_n = 1;
    for (int i = 0; i < n; i++) max_n = std::max(max_n, reqs[i].n_samples);
    const int chunks_cap = (max_n + 1023) / 1024;
    if (!s->cb || s->cb_n < n || s->cb_chunks < chunks_cap)
        {
            if (s->cb) gb200_trk_batch_destroy(s->cb);
            s->cb = nullptr;
            s->cb_n = std::max(n, s->cb_n);
            s->cb_chunks = std::max(chunks_cap, s->cb_chunks);
            int rc = trk_batch_alloc(s, s->cb_n, s->cb_chunks, &s->cb);
            if (rc)
                {
                    s->cb_n = s->cb_chunks = 0;
                    return rc;
                }
        }
    gb200_trk_batch* b = s->cb;
    b->n = n;
    const bool trace = s->tune.trace != 0;  // host wall-clock phases of one call, to stderr
    const auto t0 = std::chrono::steady_clock::now();
    int rc = trk_batch_stage(b, reqs);
    if (rc) return rc;
    const auto t1 = std::chrono::steady_clock::now();
    GB_CUDA(s, gb_copy_words(b->d_work, b->h_work.dev, sizeof(GbTrkWork) * (size_t)n, s->stream));
    s->launches++;
    rc = gb200_trk_batch_launch(b);
    const auto t2 = std::chrono::steady_clock::now();
    if (!rc) rc = gb200_trk_batch_fetch(b, outs_iq);
    if (trace)
        {
            const auto t3 = std::chrono::steady_clock::now();
            auto us = [](auto a, auto z) { return (double)std::chrono::duration_cast<std::chrono::nanoseconds>(z - a).count() * 1e-3; };
            fprintf(stderr, "gb200_trk_correlate_batch n=%d: stage %.1f us, launch %.1f us (%zu groups), wait+fetch %.1f us\n", n, us(t0, t1), us(t1, t2),
                b->groups.size(), us(t2, t3));
        }
    return rc;
}


// ------------------------------------------------------------------ int16 correlator (Cpu_Multicorrelator_16sc)
extern "C" int gb200_trk_correlate_16sc_batch(gb200_service* s, int n, const gb200_trk_req* reqs, int16_t* outs_iq)
{
    if (!s || n <= 0 || !reqs || !outs_iq) return GB200_E_INVALID;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    const size_t work_bytes = sizeof(GbTrk16Work) * (size_t)n, out_bytes = sizeof(int16_t) * 2 * GB200_MAX_TAPS * (size_t)n;
    if (s->h16_bytes < work_bytes + out_bytes)
        {
            if (s->h16) cudaFreeHost(s->h16);
            s->h16 = nullptr;
            s->h16_bytes = 0;
            GB_CUDA(s, cudaMallocHost(&s->h16, 2 * (work_bytes + out_bytes)));
            s->h16_bytes = 2 * (work_bytes + out_bytes);
        }
    GbTrk16Work* hw = (GbTrk16Work*)s->h16;
    int16_t* hout = (int16_t*)((char*)s->h16 + work_bytes);
    int max_n = 1, max_taps = 1;
    bool same_taps = true;
    for (int i = 0; i < n; i++)
        {
            const gb200_trk_req& r = reqs[i];
            gb200_trk* t = r.trk;
            if (!t || t->s != s || t->c.code_kind != GB200_CODE_16SC || t->code_len < 1 || r.n_samples < 1)
                {
                    s->set_error("gb200_trk_correlate_16sc: request " + std::to_string(i) + " needs a GB200_CODE_16SC handle with a local code and n_samples >= 1");
                    return GB200_E_INVALID;
                }
            const void* win = nullptr;
            int rc = ring_window(s, r.stream_id, r.first_sample_index, (uint64_t)r.n_samples, GB200_ITEM_16IC, &win);
            if (rc) return rc;
            GbTrk16Work& w = hw[i];
            w.samples = win;
            w.code = t->d_code;
            w.code_len = t->code_len;
            w.n = r.n_samples;
            w.n_taps = t->n_taps;
            w.out_index = i;
            memcpy(w.shifts, t->shifts, sizeof(w.shifts));
            w.rem_code = r.rem_code_phase_chips;
            w.code_step = r.code_phase_step_chips;
            w.phase0 = make_float2(cosf(r.rem_carr_phase_rad), -sinf(r.rem_carr_phase_rad));
            w.inc = make_float2(cosf(-r.phase_step_rad), sinf(-r.phase_step_rad));
            max_n = std::max(max_n, r.n_samples);
            max_taps = std::max(max_taps, t->n_taps);
            same_taps = same_taps && t->n_taps == reqs[0].trk->n_taps;
        }
    const int max_blocks = (max_n + 255) / 256;
    void *d_work, *d_ckpt, *d_triples, *d_out;
    int rc;
    if ((rc = s->scratch(10, work_bytes, &d_work))) return rc;
    if ((rc = s->scratch(11, sizeof(float2) * (size_t)n * max_blocks, &d_ckpt))) return rc;
    if ((rc = s->scratch(12, sizeof(int4) * 2 * GB200_MAX_TAPS * (size_t)n * max_blocks, &d_triples))) return rc;
    if ((rc = s->scratch(13, out_bytes, &d_out))) return rc;
    GB_CUDA(s, cudaMemcpyAsync(d_work, hw, work_bytes, cudaMemcpyHostToDevice, s->stream));
    int l = 0;
    if ((rc = rings_acquire(s))) return rc;
    GB_CUDA(s, gb_trk16_launch((const GbTrk16Work*)d_work, n, max_n, same_taps ? max_taps : GB200_MAX_TAPS, (float2*)d_ckpt, (int4*)d_triples, (short*)d_out,
                   s->stream, &l));
    if ((rc = rings_release(s))) return rc;
    s->launches += l;
    GB_CUDA(s, cudaMemcpyAsync(hout, d_out, out_bytes, cudaMemcpyDeviceToHost, s->stream));
    GB_CUDA(s, cudaStreamSynchronize(s->stream));
    memcpy(outs_iq, hout, out_bytes);
    return GB200_OK;
}

extern "C" int gb200_trk_correlate_16sc(gb200_trk* t, int stream_id, uint64_t first_sample_index, float rem_carr_phase_rad, float phase_step_rad,
    float rem_code_phase_chips, float code_phase_step_chips, int n_samples, int16_t* corr_out_iq)
{
    if (!t || !corr_out_iq) return GB200_E_INVALID;
    gb200_trk_req r;
    memset(&r, 0, sizeof(r));
    r.trk = t;
    r.stream_id = stream_id;
    r.first_sample_index = first_sample_index;
    r.rem_carr_phase_rad = rem_carr_phase_rad;
    r.phase_step_rad = phase_step_rad;
    r.rem_code_phase_chips = rem_code_phase_chips;
    r.code_phase_step_chips = code_phase_step_chips;
    r.n_samples = n_samples;
    int16_t out[2 * GB200_MAX_TAPS];
    int rc = gb200_trk_correlate_16sc_batch(t->s, 1, &r, out);
    if (rc) return rc;
    memcpy(corr_out_iq, out, sizeof(int16_t) * 2 * t->n_taps);
    return GB200_OK;
}

// one request, its own launch (the path of batch_window_us == 0, and the fall-back that isolates a failing request of a batch)
static int trk_correlate_direct(gb200_trk* t, const gb200_trk_req& r, float* corr_out_iq)
{
    gb200_service* s = t->s;
    std::lock_guard<std::recursive_mutex> lk(s->mu);
    cudaSetDevice(s->device);
    int rc;
    if (!t->single)
        {
            // per-handle resident resources: one request, up to ring_slack samples in 1024-sample chunks
            rc = trk_batch_alloc(s, 1, (int)((s->ring_slack + 1023) / 1024), &t->single);
            if (rc) return rc;
        }
    gb200_trk_batch* b = t->single;
    rc = trk_batch_stage(b, &r);
    if (rc) return rc;
    GB_CUDA(s, gb_copy_words(b->d_work, b->h_work.dev, sizeof(GbTrkWork), s->stream));
    s->launches++;
    rc = gb200_trk_batch_launch(b);
    if (rc) return rc;
    GB_CUDA(s, cudaMemcpyAsync(b->h_out, b->d_out, sizeof(float2) * GB200_OUT_STRIDE, cudaMemcpyDeviceToHost, s->stream));
    GB_CUDA(s, cudaStreamSynchronize(s->stream));
    memcpy(corr_out_iq, b->h_out, sizeof(float2) * (t->n_taps + (t->has_data ? 1 : 0)));
    return GB200_OK;
}

extern "C" int gb200_trk_correlate(gb200_trk* t, int stream_id, uint64_t first_sample_index, float rem_carr_phase_rad,
    float phase_step_rad, float phase_rate_step_rad, float rem_code_phase_chips, float code_phase_step_chips,
    float code_phase_rate_step_chips, int n_samples, float* corr_out_iq)
{
    if (!t || !corr_out_iq) return GB200_E_INVALID;
    gb200_service* s = t->s;
    gb200_trk_req r;
    r.trk = t;
    r.stream_id = stream_id;
    r.first_sample_index = first_sample_index;
    r.rem_carr_phase_rad = rem_carr_phase_rad;
    r.phase_step_rad = phase_step_rad;
    r.phase_rate_step_rad = phase_rate_step_rad;
    r.rem_code_phase_chips = rem_code_phase_chips;
    r.code_phase_step_chips = code_phase_step_chips;
    r.code_phase_rate_step_chips = code_phase_rate_step_chips;
    r.n_samples = n_samples;
    if (s->batch_window_us == 0 || t->c.code_kind == GB200_CODE_16SC) return trk_correlate_direct(t, r, corr_out_iq);

    // ---- coalescing: every channel's thread queues its epoch; the first one in collects for at most batch_window_us (or until
    // every live channel has arrived), launches ONE batch for all of them and hands the results back ----
    gb200_service::TrkPending me{r, corr_out_iq, GB200_OK, false};
    std::unique_lock<std::mutex> lk(s->qmu);
    s->queue.push_back(&me);
    if (s->collecting)
        {
            s->qcv.notify_all();  // the collector re-checks "everyone here?"
            s->qcv.wait(lk, [&] { return me.done; });
            return me.rc;
        }
    s->collecting = true;
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::microseconds(s->batch_window_us);
    s->qcv.wait_until(lk, deadline, [&] {
        const int have = (int)s->queue.size();
        return have >= s->live_trk || (!s->running && have >= std::max(1, std::min(s->prev_batch_n, s->live_trk)));
    });
    // the GPU lane is still busy with the previous batch: nothing is lost by collecting until it is done
    s->qcv.wait(lk, [&] { return !s->running; });
    std::vector<gb200_service::TrkPending*> batch;
    batch.swap(s->queue);
    s->prev_batch_n = (int)batch.size();
    s->running = true;
    s->collecting = false;  // the next arrival starts collecting the next batch while this one runs
    lk.unlock();
    const int n = (int)batch.size();
    std::vector<gb200_trk_req> reqs(n);
    for (int i = 0; i < n; i++) reqs[i] = batch[i]->req;
    std::vector<float> outs((size_t)n * 2 * GB200_OUT_STRIDE);
    int rc = gb200_trk_correlate_batch(s, n, reqs.data(), outs.data());
    if (rc != GB200_OK && n > 1)
        {
            // one bad request (e.g. samples already overwritten in the ring) must not fail the others: redo one by one
            for (int i = 0; i < n; i++) batch[i]->rc = trk_correlate_direct(batch[i]->req.trk, batch[i]->req, batch[i]->out);
        }
    else
        for (int i = 0; i < n; i++)
            {
                batch[i]->rc = rc;
                const gb200_trk* ti = batch[i]->req.trk;
                if (rc == GB200_OK) memcpy(batch[i]->out, &outs[(size_t)i * 2 * GB200_OUT_STRIDE], sizeof(float2) * (ti->n_taps + (ti->has_data ? 1 : 0)));
            }
    lk.lock();
    s->coalesced_batches++;
    s->coalesced_requests += (uint64_t)n;
    for (int i = 0; i < n; i++) batch[i]->done = true;
    s->running = false;
    const int my_rc = me.rc;
    lk.unlock();
    s->qcv.notify_all();
    return my_rc;
}

extern "C" int gb200_service_coalescing_stats(gb200_service* s, uint64_t* batches, uint64_t* requests)
{
    if (!s) return GB200_E_INVALID;
    std::lock_guard<std::mutex> lk(s->qmu);
    if (batches) *batches = s->coalesced_batches;
    if (requests) *requests = s->coalesced_requests;
    return GB200_OK;
}
