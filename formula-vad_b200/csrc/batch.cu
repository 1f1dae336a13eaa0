// batch.cu -- the AudioPipeline drop-in: a batch of independent pipelines on one GPU.
// Host-side orchestration of src/AudioPipeline.zig (pushSamples) and src/AudioPipeline/VAD.zig
// (collectInputStep -> preAnalyze -> denoiserStep -> fftBufferStep -> fftStep -> stateMachineStep),
// re-expressed as whole-tile kernel launches over (streams x frames).
#include "common.cuh"
#include "frames.cuh"
#include "rnnoise.cuh"
#include "stft.cuh"
#include "vadm.cuh"
#include <algorithm>
#include <atomic>
#include <cmath>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <new>
#include <vector>

using namespace fvad;

// Pageable host input (what a Zig allocator hands over): cudaMemcpyAsync from pageable memory is a staged, synchronous
// copy at ~10 GB/s -- 0.43 of the device-resident rate at 1184 streams.  The batch therefore owns a small ring of PINNED
// staging slots and a few copy threads: the rows of a chunk are copied into a slot by the threads while the DMA engine
// drains the previous slot, and the push returns as soon as its samples have left the caller's memory.
class HostCopyPool {
public:
    explicit HostCopyPool(int n_threads) {
        for (int i = 0; i < n_threads; i++) workers_.emplace_back([this] { run(); });
    }
    ~HostCopyPool() {
        { std::lock_guard<std::mutex> lk(mu_); stop_ = true; }
        cv_.notify_all();
        for (auto &t : workers_) t.join();
    }
    // fn(i) for i in [0, n), spread over the threads; returns when all are done
    void parallel_for(size_t n, const std::function<void(size_t)> &fn) {
        if (n == 0) return;
        {
            std::lock_guard<std::mutex> lk(mu_);
            fn_ = &fn; next_ = 0; n_ = n; pending_ = n; gen_++;
        }
        cv_.notify_all();
        std::unique_lock<std::mutex> lk(mu_);
        done_cv_.wait(lk, [this] { return pending_ == 0; });
        fn_ = nullptr;
    }
private:
    void run() {
        unsigned long long seen = 0;
        for (;;) {
            std::unique_lock<std::mutex> lk(mu_);
            cv_.wait(lk, [&] { return stop_ || (gen_ != seen && next_ < n_); });
            if (stop_) return;
            while (next_ < n_) {
                const size_t i = next_++;
                const std::function<void(size_t)> *fn = fn_;
                lk.unlock();
                (*fn)(i);
                lk.lock();
                if (--pending_ == 0) done_cv_.notify_all();
            }
            seen = gen_;
        }
    }
    std::vector<std::thread> workers_;
    std::mutex mu_;
    std::condition_variable cv_, done_cv_;
    const std::function<void(size_t)> *fn_ = nullptr;
    size_t next_ = 0, n_ = 0, pending_ = 0;
    unsigned long long gen_ = 0;
    bool stop_ = false;
};

struct fvad_batch {
    fvad_config cfg;
    std::vector<fvad_vad_machine_config> machine_cfgs;   // [0] main, [1..] alt
    size_t S = 0, C = 0, rows = 0;
    int device = 0;
    // Six stage streams (wavefront over tiles): sA biquad, sB analysis/pitch, sB2 pitch decision scan + features, sC RNN,
    // stream (= stage D) synthesis + STFT, sE VADMachine.  The latency-bound scans (pitch decision, VADMachine: one thread
    // per stream walking the tile's frames) have stages of their own so that the frame-parallel kernels of the next tile
    // run beside them instead of behind them.  `stream` is the batch's main stream: it joins sE at the end of every push,
    // so results become visible in its order.
    cudaStream_t stream = nullptr, sA = nullptr, sB = nullptr, sB2 = nullptr, sC = nullptr, sE = nullptr;
    static constexpr int kMaxSlots = 8;
    int n_slots = 1;
    cudaEvent_t evA[kMaxSlots] = {}, evB1[kMaxSlots] = {}, evB[kMaxSlots] = {}, evC[kMaxSlots] = {}, evD[kMaxSlots] = {};
    cudaEvent_t evS[2] = {}, evE[2] = {};   // per parity of the STFT-rate buffers: produced (stage D) / consumed (stage E)
    unsigned long long stft_tile_no = 0;    // tiles that reached the machines: parity = stft_tile_no & 1
    int e_pending = -1;                     // parity whose evE the main stream has not joined yet
    unsigned long long tile_no = 0;
    float *d_ratio_slot[kMaxSlots] = {};   // [S][tile_frames] channel-volume ratio per 480-frame, stage A -> D
    bool use_denoiser = true;
    int fft_size = 2048, frame_len = 480;
    unsigned trace_flags = 0;

    // tiling
    size_t tile_frames = 0;        // frames of frame_len per row per tile
    size_t work_cap = 0;           // floats per work row
    // Double-buffered work rows [rows][work_cap]: the next tile's PCM is copied (stream sH) into one buffer while stage A
    // reads the other, so the copy of tile t+1 hides behind the biquad chain of tile t.
    float *d_work_buf[2] = {nullptr, nullptr};
    // Pushes APPEND to the filling buffer (copy stream sH: one 2-D copy per push when the caller's rows are equally
    // strided); its tile is launched when it holds tile_frames frames, or at a sync.  So the launch count depends on the
    // audio pushed, not on the caller's push size (SimulationInstance.simulateVAD pushes one second at a time), and the
    // copy of tile t+1 hides behind stage A of tile t.
    int work_par = 0;              // the filling buffer
    size_t fill_len = 0;           // valid samples per row in it (a carried partial frame included)
    bool fill_waited = false;      // sH already waits for the buffer's previous reader
    bool h2d_dirty = false;        // copies issued on sH since the last evH record
    cudaStream_t sH = nullptr;
    cudaEvent_t evH[2] = {}, evW[2] = {};   // per work buffer: input copied in (sH) / last reader done (input-side stream)

    // denoiser path
    RnnoiseBatch *rn = nullptr;
    float *d_den = nullptr;        // [rows][den_cap]: carried denoised samples + the tile's output
    size_t den_cap = 0, den_carry = 0;
    float *d_den_carry = nullptr;  // [rows][fft_size]
    static constexpr int kHist480 = 16;         // 480-sample frames of history in front of a tile: an STFT frame reaches fft_size/480 + 2 back
    float *d_ratio480 = nullptr, *d_vad480 = nullptr;   // [S][kHist480 + tile_frames]
    size_t stride480 = 0;
    long long base480 = 0;         // global 480-frame number of slot kHist480

    // STFT + machines
    const StftTables *tabs = nullptr;
    StftBands bands{};
    size_t max_stft = 0;           // STFT frames per row per tile
    float *d_band = nullptr;       // [2][n_bands][rows][max_stft]   (two parities: stage D writes one while stage E reads the other)
    float *d_ratio_stft = nullptr, *d_vad_stft = nullptr;   // [2][S][max_stft]
    float *d_spectra = nullptr;    // trace only: [rows][max_stft][1025]
    int n_machines = 0;
    VadmParams *d_params = nullptr;
    VadmState *d_states = nullptr;
    double *d_rings = nullptr;
    std::vector<VadmParams> h_params;
    size_t rings_per_stream = 0;
    // result arenas: appended on the device, drained at a sync -- or earlier, when the counters (mirrored into pinned host
    // memory after every tile, read without waiting) say an arena is half full
    unsigned *h_arena_cnt = nullptr;       // pinned [2]: segments, events (a value the device has reached: a lower bound)
    // FVAD_KEEP_FEATURES: the per-STFT-frame features of the whole run stay in HBM for fvad_batch_sweep
    float *d_feat_min = nullptr, *d_feat_ratio = nullptr, *d_feat_vad = nullptr;   // [bands][S][feat_cap], [S][feat_cap] x 2
    size_t feat_cap = 0;
    // pageable host input: pinned staging slots + copy threads (created at the first pageable push)
    static constexpr int kStageSlots = 3;
    float *h_stage[kStageSlots] = {};      // pinned, [rows][stage_samples] each
    cudaEvent_t evStage[kStageSlots] = {}; // the H2D copy out of the slot has completed
    size_t stage_samples = 0;
    int stage_next = 0;
    HostCopyPool *copy_pool = nullptr;
    // Recorder (fvad_batch_set_recording): history of the original PCM + the host-side capture state per stream
    bool rec_on = false;
    bool hist_fixed = false;               // a ring of hist_cap samples (else: everything pushed, growing)
    float *d_hist = nullptr;               // [rows][hist_cap]: absolute sample i of a row at [i % hist_cap]
    size_t hist_cap = 0;
    uint64_t hist_len = 0;                 // samples per row copied into the history so far
    fvad_on_recording_fn on_recording = nullptr;
    void *on_recording_ctx = nullptr;
    struct RecState { bool active = false; uint64_t start = 0; size_t next_event = 0; };
    struct PendingClip { size_t stream; uint64_t start, end; };
    struct StoredClip { fvad_clip meta; std::vector<float> pcm; };
    std::vector<RecState> rec_state;       // [stream]
    std::vector<PendingClip> rec_pending;  // completed by the machine, waiting for the stream to reach their end
    std::vector<std::vector<StoredClip>> rec_clips;     // [stream] delivered since the last fvad_batch_clips
    std::vector<std::vector<fvad_clip>> rec_view;       // [stream] what the last fvad_batch_clips returned
    std::vector<std::vector<StoredClip>> rec_held;      // [stream] storage behind rec_view
    ClipPiece *d_clip_pieces = nullptr; size_t clip_pieces_cap = 0;
    float *d_clip_stage = nullptr; size_t clip_stage_cap = 0;
    // sweep scratch, kept between calls
    VadmParams *d_sw_params = nullptr; size_t sw_params_cap = 0;
    double *d_sw_rings = nullptr; size_t sw_rings_cap = 0;
    fvad_segment *d_sw_segs = nullptr; size_t sw_segs_cap = 0;
    uint32_t *d_sw_counts = nullptr; size_t sw_counts_cap = 0;
    SegRecord *d_segs = nullptr; unsigned *d_seg_count = nullptr; unsigned seg_cap = 0;
    EventRecord *d_events = nullptr; unsigned *d_event_count = nullptr; unsigned event_cap = 0;
    fvad_stft_record *d_trace = nullptr;   // [S][max_stft]

    // progress
    uint64_t total_write_count = 0, read_count = 0, stft_done = 0;

    // results (host)
    std::vector<std::vector<fvad_segment>> segs;        // [stream*n_machines + k]
    std::vector<std::vector<fvad_capture_event>> events;   // [stream]
    std::vector<std::vector<fvad_stft_record>> tr_records; // [stream]
    std::vector<std::vector<float>> tr_denoised;        // [row]
    std::vector<std::vector<float>> tr_vad;             // [stream]
    std::vector<std::vector<float>> tr_spectra;         // [row]
    bool dirty = false;
};

namespace {

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

template <class T> int dmalloc(T **p, size_t n) {
    cudaError_t e = cudaMalloc((void **)p, n * sizeof(T) + 16);
    if (e != cudaSuccess) {
        *p = nullptr;
        return set_error(e == cudaErrorMemoryAllocation ? FVAD_EOUT_OF_MEMORY : FVAD_ECUDA, "cudaMalloc(%zu bytes): %s",
                         n * sizeof(T), cudaGetErrorString(e));
    }
    return FVAD_OK;
}

int freq_to_bin(size_t sample_rate, size_t n_fft, float freq, int *bin) {   // FFT.zig:120-131
    const float nyq = (float)sample_rate / 2;
    if (freq > nyq) return set_error(FVAD_EOUT_OF_RANGE, "frequency %g above Nyquist", (double)freq);
    if (freq < 0) return set_error(FVAD_ENEGATIVE_FREQUENCY, "negative frequency %g", (double)freq);
    const float bw = (float)sample_rate / (float)n_fft;
    *bin = (int)roundf(freq / bw);
    return FVAD_OK;
}

int upload_machine_state(fvad_batch *b) {       // VADMachine.init for every (stream, machine), on the device
    FVAD_TRY(vadm_init_launch(b->d_params, b->d_states, b->d_rings, b->S, b->n_machines, b->stream));
    FVAD_CUDA_TRY(cudaMemsetAsync(b->d_seg_count, 0, sizeof(unsigned), b->stream));
    FVAD_CUDA_TRY(cudaMemsetAsync(b->d_event_count, 0, sizeof(unsigned), b->stream));
    if (b->h_arena_cnt) b->h_arena_cnt[0] = b->h_arena_cnt[1] = 0;
    return FVAD_OK;
}

}  // namespace

extern "C" void fvad_batch_destroy(fvad_batch *b) {
    if (!b) return;
    {
        DeviceGuard g(b->device);
        for (cudaStream_t st : {b->sA, b->sB, b->sB2, b->sC, b->sE, b->stream}) if (st) cudaStreamSynchronize(st);
        if (b->rn) rnnoise_batch_destroy(b->rn);
        for (int i = 0; i < fvad_batch::kMaxSlots; i++) {
            cudaFree(b->d_ratio_slot[i]);
            for (cudaEvent_t e : {b->evA[i], b->evB1[i], b->evB[i], b->evC[i], b->evD[i]}) if (e) cudaEventDestroy(e);
        }
        for (cudaEvent_t e : {b->evS[0], b->evS[1], b->evE[0], b->evE[1]}) if (e) cudaEventDestroy(e);
        if (b->sH) { cudaStreamSynchronize(b->sH); cudaStreamDestroy(b->sH); }
        for (cudaEvent_t e : {b->evH[0], b->evH[1], b->evW[0], b->evW[1]}) if (e) cudaEventDestroy(e);
        cudaFree(b->d_work_buf[0]); cudaFree(b->d_work_buf[1]); cudaFree(b->d_den); cudaFree(b->d_den_carry);
        cudaFree(b->d_ratio480); cudaFree(b->d_vad480); cudaFree(b->d_band); cudaFree(b->d_ratio_stft);
        cudaFree(b->d_vad_stft); cudaFree(b->d_spectra); cudaFree(b->d_params); cudaFree(b->d_states);
        cudaFree(b->d_rings); cudaFree(b->d_segs); cudaFree(b->d_seg_count); cudaFree(b->d_events);
        cudaFree(b->d_event_count); cudaFree(b->d_trace);
        cudaFree(b->d_feat_min); cudaFree(b->d_feat_ratio); cudaFree(b->d_feat_vad);
        cudaFree(b->d_hist); cudaFree(b->d_clip_pieces); cudaFree(b->d_clip_stage);
        cudaFree(b->d_sw_params); cudaFree(b->d_sw_rings); cudaFree(b->d_sw_segs); cudaFree(b->d_sw_counts);
        if (b->h_arena_cnt) cudaFreeHost(b->h_arena_cnt);
        delete b->copy_pool;
        for (int i = 0; i < fvad_batch::kStageSlots; i++) {
            if (b->h_stage[i]) cudaFreeHost(b->h_stage[i]);
            if (b->evStage[i]) cudaEventDestroy(b->evStage[i]);
        }
        for (cudaStream_t st : {b->sA, b->sB, b->sB2, b->sC, b->sE}) if (st && st != b->stream) cudaStreamDestroy(st);
        if (b->stream) cudaStreamDestroy(b->stream);
    }
    delete b;
}

extern "C" int fvad_batch_create(const fvad_config *cfg, size_t n_streams, int device, fvad_batch **out) {
    if (!cfg || !out || n_streams == 0 || cfg->n_channels == 0)
        return set_error(FVAD_EINVALID_ARGUMENT, "fvad_batch_create: null/zero argument");
    *out = nullptr;
    if (cfg->sample_rate != 48000)                                             // VAD.zig:101-104
        return set_error(FVAD_EINVALID_SAMPLE_RATE, "RNNoise can only handle 48 kHz audio (got %zu)", cfg->sample_rate);
    const size_t n_fft = cfg->vad_config.fft_size;
    if (n_fft == 0 || (n_fft % 2) != 0)                                        // FFT.zig:29-31
        return set_error(FVAD_EINVALID_FFT_SIZE, "fft_size %zu must be even and non-zero", n_fft);
    if (n_fft != (size_t)kFft) {                                               // any other size: the generic plan (stft_generic.cu)
        const int rc_fft = stft_generic_supported(n_fft);
        if (rc_fft != FVAD_OK) return rc_fft;
    }
    if (cfg->n_channels > 64) return set_error(FVAD_EUNSUPPORTED, "more than 64 channels");
    if (cfg->vad_config.use_denoiser && n_fft / kRnFrame + 3 > (size_t)fvad_batch::kHist480)
        return set_error(FVAD_EUNSUPPORTED, "fft_size %zu with the denoiser: the re-framing history covers sizes up to %d", n_fft,
                         (fvad_batch::kHist480 - 3) * kRnFrame);
    if (device < 0) {
        cudaError_t e = cudaGetDevice(&device);
        if (e != cudaSuccess) return set_error(FVAD_ECUDA, "cudaGetDevice: %s (no CPU fallback exists)", cudaGetErrorString(e));
    }
    DeviceGuard g(device);
    if (!g.ok) return set_error(FVAD_ECUDA, "cannot select CUDA device %d (no CPU fallback exists)", device);

    fvad_batch *b = new (std::nothrow) fvad_batch();
    if (!b) return set_error(FVAD_EOUT_OF_MEMORY, "host allocation failed");
    b->cfg = *cfg;
    b->S = n_streams; b->C = cfg->n_channels; b->rows = b->S * b->C; b->device = device;
    b->use_denoiser = cfg->vad_config.use_denoiser != 0;
    b->fft_size = (int)n_fft;
    b->frame_len = b->use_denoiser ? kRnFrame : b->fft_size;                    // VAD.zig:206-212
    b->machine_cfgs.push_back(cfg->vad_config.vad_machine_config);
    for (size_t i = 0; i < cfg->vad_config.n_alt_vad_machine_configs; i++)
        b->machine_cfgs.push_back(cfg->vad_config.alt_vad_machine_configs[i]);
    b->cfg.vad_config.alt_vad_machine_configs = nullptr;                        // we own a copy
    b->n_machines = (int)b->machine_cfgs.size();

    int rc = FVAD_OK;
    auto fail = [&](int code) { fvad_batch_destroy(b); return code; };
    {
        // Stream priorities decide whose CTAs the block scheduler places first when SM resources free up; rank 0 = most urgent.
        // Latency regime (few chains): stage A carries the chain that sets the pace, so it goes first -- without that the
        // tile's small stage-A kernels queue behind the ten-thousands of CTAs stage B of the previous tile has just launched
        // (measured: ~5 ms per tile, 6 % of the step); then C, then B, then D in chain-length order.
        // Throughput regime (SMs saturated): what matters is that the SMALL kernels between the big ones (history copies,
        // scans, carries: a few CTAs each) do not wait for a 30 000-CTA grid to be dispatched -- measured with
        // knob timeline: a 10 us copy took 3 ms -- and that older tiles drain first, so later stages outrank earlier ones.
        // knob prio="a,b1,b2,c,d,e" overrides the ranks.
        int lo = 0, hi = 0;
        if (cudaDeviceGetStreamPriorityRange(&lo, &hi) != cudaSuccess) lo = hi = 0;      // hi = numerically lowest = most urgent
        int rank[6] = {0, 2, 2, 1, 99, 2};                                               // A, B1, B2, C, D, E
        if (!latency_regime(n_streams, b->device)) { const int thr[6] = {0, 5, 4, 3, 2, 1}; for (int i = 0; i < 6; i++) rank[i] = thr[i]; }
        if (const char *e = debug_knob("prio")) {
            int v[6];
            if (sscanf(e, "%d,%d,%d,%d,%d,%d", &v[0], &v[1], &v[2], &v[3], &v[4], &v[5]) == 6) for (int i = 0; i < 6; i++) rank[i] = v[i];
        }
        auto prio = [&](int r) { const int p = hi + (r < 0 ? 0 : r); return p > lo ? lo : p; };
        if (cudaStreamCreateWithPriority(&b->stream, cudaStreamNonBlocking, prio(rank[4])) != cudaSuccess)
            return fail(set_error(FVAD_ECUDA, "cudaStreamCreate failed: %s", cudaGetErrorString(cudaGetLastError())));
        if (debug_knob("serial_stages")) {
            // measurement aid: all stages on one stream, so the per-kernel event times are free of overlap
            b->sA = b->sB = b->sB2 = b->sC = b->sE = b->stream;
        } else if (cudaStreamCreateWithPriority(&b->sA, cudaStreamNonBlocking, prio(rank[0])) != cudaSuccess ||
                   cudaStreamCreateWithPriority(&b->sB, cudaStreamNonBlocking, prio(rank[1])) != cudaSuccess ||
                   cudaStreamCreateWithPriority(&b->sB2, cudaStreamNonBlocking, prio(rank[2])) != cudaSuccess ||
                   cudaStreamCreateWithPriority(&b->sC, cudaStreamNonBlocking, prio(rank[3])) != cudaSuccess ||
                   cudaStreamCreateWithPriority(&b->sE, cudaStreamNonBlocking, prio(rank[5])) != cudaSuccess)
            return fail(set_error(FVAD_ECUDA, "cudaStreamCreate failed: %s", cudaGetErrorString(cudaGetLastError())));
    }
    {
        const char *es = debug_knob("slots");
        int ns = es ? atoi(es) : 3;
        b->n_slots = ns < 1 ? 1 : ns > fvad_batch::kMaxSlots ? fvad_batch::kMaxSlots : ns;
        for (int i = 0; i < b->n_slots; i++)
            for (cudaEvent_t *e : {&b->evA[i], &b->evB1[i], &b->evB[i], &b->evC[i], &b->evD[i]})
                if (cudaEventCreateWithFlags(e, cudaEventDisableTiming) != cudaSuccess)
                    return fail(set_error(FVAD_ECUDA, "cudaEventCreate failed: %s", cudaGetErrorString(cudaGetLastError())));
        for (cudaEvent_t *e : {&b->evS[0], &b->evS[1], &b->evE[0], &b->evE[1]})
            if (cudaEventCreateWithFlags(e, cudaEventDisableTiming) != cudaSuccess)
                return fail(set_error(FVAD_ECUDA, "cudaEventCreate failed: %s", cudaGetErrorString(cudaGetLastError())));
    }
    if ((rc = stft_tables_get(device, &b->tabs)) != FVAD_OK) return fail(rc);

    // machines: params, distinct bands (PipelineFFT.averageVolumeInBand per machine config)
    b->h_params.resize(b->n_machines);
    size_t ring_off = 0;
    for (int k = 0; k < b->n_machines; k++) {
        size_t nd = 0;
        if ((rc = vadm_params_from_config(b->machine_cfgs[k], cfg->sample_rate, n_fft, &b->h_params[k], &nd)) != FVAD_OK) return fail(rc);
        int lo = 0, hi = 0;
        if ((rc = freq_to_bin(cfg->sample_rate, n_fft, b->machine_cfgs[k].speech_min_freq, &lo)) != FVAD_OK) return fail(rc);
        if ((rc = freq_to_bin(cfg->sample_rate, n_fft, b->machine_cfgs[k].speech_max_freq, &hi)) != FVAD_OK) return fail(rc);
        int band = -1;
        for (int i = 0; i < b->bands.n_bands; i++) if (b->bands.lo[i] == lo && b->bands.hi[i] == hi) band = i;
        if (band < 0) {
            if (b->bands.n_bands == kMaxBands) return fail(set_error(FVAD_EUNSUPPORTED, "more than %d distinct speech bands", kMaxBands));
            band = b->bands.n_bands++;
            b->bands.lo[band] = lo; b->bands.hi[band] = hi;
        }
        b->h_params[k].band = band;
        b->h_params[k].ring_off = ring_off;
        ring_off += nd;
    }
    b->rings_per_stream = ring_off;

    // tile geometry
    const char *env = debug_knob("tile_frames");
    size_t tf;
    if (b->use_denoiser) {
        tf = 131072 / b->rows;
        if (tf > 6000) tf = 6000;
        if (tf < 64) tf = 64;
    } else {
        tf = ((size_t)1 << 28) / (b->rows * (size_t)b->fft_size);              // ~1 GiB of f32 per tile
        if (tf > 4096) tf = 4096;
        if (tf < 4) tf = 4;
    }
    if (env && atol(env) > 0) tf = (size_t)atol(env);
    b->tile_frames = tf;
    b->work_cap = tf * b->frame_len + b->frame_len;
    b->work_cap = (b->work_cap + 3) & ~(size_t)3;
    for (int i = 0; i < 2; i++)
        if ((rc = dmalloc(&b->d_work_buf[i], b->rows * b->work_cap)) != FVAD_OK) return fail(rc);
    if (cudaStreamCreateWithFlags(&b->sH, cudaStreamNonBlocking) != cudaSuccess)
        return fail(set_error(FVAD_ECUDA, "cudaStreamCreate failed: %s", cudaGetErrorString(cudaGetLastError())));
    for (cudaEvent_t *e : {&b->evH[0], &b->evH[1], &b->evW[0], &b->evW[1]})
        if (cudaEventCreateWithFlags(e, cudaEventDisableTiming) != cudaSuccess)
            return fail(set_error(FVAD_ECUDA, "cudaEventCreate failed: %s", cudaGetErrorString(cudaGetLastError())));
    if (b->use_denoiser) {
        b->den_cap = (tf * kRnFrame + b->fft_size + 3) & ~(size_t)3;
        b->stride480 = fvad_batch::kHist480 + tf;
        if ((rc = dmalloc(&b->d_den, b->rows * b->den_cap)) != FVAD_OK) return fail(rc);
        if ((rc = dmalloc(&b->d_den_carry, b->rows * (size_t)b->fft_size)) != FVAD_OK) return fail(rc);
        if ((rc = dmalloc(&b->d_ratio480, b->S * b->stride480)) != FVAD_OK) return fail(rc);
        if ((rc = dmalloc(&b->d_vad480, b->S * b->stride480)) != FVAD_OK) return fail(rc);
        if ((rc = rnnoise_batch_create(b->S, tf * b->C, b->n_slots, device, &b->rn)) != FVAD_OK) return fail(rc);
        for (int i = 0; i < b->n_slots; i++)
            if ((rc = dmalloc(&b->d_ratio_slot[i], b->S * tf)) != FVAD_OK) return fail(rc);
        b->max_stft = (tf * kRnFrame + b->fft_size) / b->fft_size + 1;
    } else {
        b->max_stft = tf;
    }
    if ((rc = dmalloc(&b->d_band, 2 * (size_t)b->bands.n_bands * b->rows * b->max_stft)) != FVAD_OK) return fail(rc);
    if ((rc = dmalloc(&b->d_ratio_stft, 2 * b->S * b->max_stft)) != FVAD_OK) return fail(rc);
    if ((rc = dmalloc(&b->d_vad_stft, 2 * b->S * b->max_stft)) != FVAD_OK) return fail(rc);
    if ((rc = dmalloc(&b->d_params, (size_t)b->n_machines)) != FVAD_OK) return fail(rc);
    if ((rc = dmalloc(&b->d_states, b->S * (size_t)b->n_machines)) != FVAD_OK) return fail(rc);
    if ((rc = dmalloc(&b->d_rings, b->S * b->rings_per_stream)) != FVAD_OK) return fail(rc);
    {
        // A machine completes at most one segment (and reports at most two state changes) per three STFT frames; up to
        // n_slots + 2 tiles are in flight behind the mirrored counters.  Four times that, within [2^20, 2^24] records.
        const size_t per_tile = b->S * (size_t)b->n_machines * (b->max_stft / 3 + 1);
        const size_t want = 4 * (size_t)(b->n_slots + 2) * per_tile;
        b->seg_cap = (unsigned)std::min<size_t>(std::max<size_t>(want, (size_t)1 << 20), (size_t)1 << 24);
        b->event_cap = (unsigned)std::min<size_t>(std::max<size_t>(2 * want / (size_t)b->n_machines, (size_t)1 << 20), (size_t)1 << 24);
    }
    if (cudaHostAlloc((void **)&b->h_arena_cnt, 2 * sizeof(unsigned), cudaHostAllocDefault) != cudaSuccess)
        return fail(set_error(FVAD_EOUT_OF_MEMORY, "cudaHostAlloc failed: %s", cudaGetErrorString(cudaGetLastError())));
    b->h_arena_cnt[0] = b->h_arena_cnt[1] = 0;
    if ((rc = dmalloc(&b->d_segs, (size_t)b->seg_cap)) != FVAD_OK) return fail(rc);
    if ((rc = dmalloc(&b->d_events, (size_t)b->event_cap)) != FVAD_OK) return fail(rc);
    if ((rc = dmalloc(&b->d_seg_count, (size_t)1)) != FVAD_OK) return fail(rc);
    if ((rc = dmalloc(&b->d_event_count, (size_t)1)) != FVAD_OK) return fail(rc);
    if (cudaMemcpyAsync(b->d_params, b->h_params.data(), sizeof(VadmParams) * b->n_machines, cudaMemcpyHostToDevice, b->stream) != cudaSuccess)
        return fail(set_error(FVAD_ECUDA, "upload of machine params failed"));
    if ((rc = upload_machine_state(b)) != FVAD_OK) return fail(rc);
    if (b->use_denoiser) {
        if (cudaMemsetAsync(b->d_ratio480, 0, sizeof(float) * b->S * b->stride480, b->stream) != cudaSuccess ||
            cudaMemsetAsync(b->d_vad480, 0, sizeof(float) * b->S * b->stride480, b->stream) != cudaSuccess)
            return fail(set_error(FVAD_ECUDA, "memset failed"));
    }
    b->segs.resize(b->S * b->n_machines);
    b->events.resize(b->S);
    b->tr_records.resize(b->S); b->tr_vad.resize(b->S);
    b->tr_denoised.resize(b->rows); b->tr_spectra.resize(b->rows);
    if (cudaStreamSynchronize(b->stream) != cudaSuccess)
        return fail(set_error(FVAD_ECUDA, "stream sync after create failed: %s", cudaGetErrorString(cudaGetLastError())));
    *out = b;
    return FVAD_OK;
}

// ------------------------------------------------------------------------------------------
// ------------------------------------------------------------------------------------------
constexpr unsigned kHostTraces = FVAD_TRACE_RECORDS | FVAD_TRACE_DENOISED | FVAD_TRACE_RNN_VAD | FVAD_TRACE_SPECTRA;
static int trace_tile(fvad_batch *b, size_t F, size_t J, size_t den_before) {
    FVAD_CUDA_TRY(cudaStreamSynchronize(b->stream));
    if (b->sE) FVAD_CUDA_TRY(cudaStreamSynchronize(b->sE));   // the records trace is written by the machines
    if ((b->trace_flags & FVAD_TRACE_RECORDS) && J) {
        std::vector<fvad_stft_record> tmp(J);
        for (size_t s = 0; s < b->S; s++) {
            FVAD_CUDA_TRY(cudaMemcpy(tmp.data(), b->d_trace + s * b->max_stft, sizeof(fvad_stft_record) * J, cudaMemcpyDeviceToHost));
            b->tr_records[s].insert(b->tr_records[s].end(), tmp.begin(), tmp.end());
        }
    }
    if ((b->trace_flags & FVAD_TRACE_SPECTRA) && J) {
        const size_t bins = (size_t)b->fft_size / 2 + 1;
        std::vector<float> tmp(J * bins);
        for (size_t r = 0; r < b->rows; r++) {
            FVAD_CUDA_TRY(cudaMemcpy(tmp.data(), b->d_spectra + r * J * bins, sizeof(float) * J * bins, cudaMemcpyDeviceToHost));
            b->tr_spectra[r].insert(b->tr_spectra[r].end(), tmp.begin(), tmp.end());
        }
    }
    if (b->use_denoiser && F) {
        if (b->trace_flags & FVAD_TRACE_DENOISED) {
            std::vector<float> tmp(F * kRnFrame);
            for (size_t r = 0; r < b->rows; r++) {
                FVAD_CUDA_TRY(cudaMemcpy(tmp.data(), b->d_den + r * b->den_cap + den_before, sizeof(float) * tmp.size(), cudaMemcpyDeviceToHost));
                b->tr_denoised[r].insert(b->tr_denoised[r].end(), tmp.begin(), tmp.end());
            }
        }
        if (b->trace_flags & FVAD_TRACE_RNN_VAD) {
            std::vector<float> tmp(F);
            for (size_t s = 0; s < b->S; s++) {
                FVAD_CUDA_TRY(cudaMemcpy(tmp.data(), b->d_vad480 + s * b->stride480 + fvad_batch::kHist480, sizeof(float) * F, cudaMemcpyDeviceToHost));
                b->tr_vad[s].insert(b->tr_vad[s].end(), tmp.begin(), tmp.end());
            }
        }
    }
    return FVAD_OK;
}

// PipelineFFT.fft for the batch's fft_size: the specialised kernel for 2048, KissFFT's generic plan otherwise
static int batch_stft(fvad_batch *b, const float *in, size_t in_stride, size_t n_frames, float *d_band, float *spec, cudaStream_t st) {
    if (b->fft_size == kFft)
        return stft_launch(*b->tabs, in, b->rows, in_stride, n_frames, b->fft_size, b->bands, d_band, b->max_stft, 0, spec, nullptr, 0.f,
                           b->device, st);
    return stft_generic_launch(b->device, (size_t)b->fft_size, in, b->rows, in_stride, n_frames, b->fft_size, b->bands, d_band, b->max_stft,
                               0, spec, nullptr, 0.f, st);
}

static float *band_buf(fvad_batch *b, int par) { return b->d_band + (size_t)par * b->bands.n_bands * b->rows * b->max_stft; }
static float *ratio_buf(fvad_batch *b, int par) { return b->d_ratio_stft + (size_t)par * b->S * b->max_stft; }
static float *vad_buf(fvad_batch *b, int par) { return b->d_vad_stft + (size_t)par * b->S * b->max_stft; }

// The kept features grow with the run: capacity doubles (a rare stop-the-world copy; every stream of the batch waits).
static int ensure_feature_capacity(fvad_batch *b, size_t need) {
    if (need <= b->feat_cap) return FVAD_OK;
    size_t cap = b->feat_cap ? b->feat_cap : 4096;
    while (cap < need) cap *= 2;
    for (cudaStream_t st : {b->sH, b->sA, b->sB, b->sB2, b->sC, b->sE, b->stream}) if (st) FVAD_CUDA_TRY(cudaStreamSynchronize(st));
    float *nm = nullptr, *nr = nullptr, *nv = nullptr;
    const size_t rows_min = (size_t)b->bands.n_bands * b->S;
    FVAD_TRY(dmalloc(&nm, rows_min * cap));
    FVAD_TRY(dmalloc(&nr, b->S * cap));
    FVAD_TRY(dmalloc(&nv, b->S * cap));
    if (b->feat_cap && b->stft_done) {
        const size_t w = b->stft_done * sizeof(float);
        FVAD_CUDA_TRY(cudaMemcpy2D(nm, cap * sizeof(float), b->d_feat_min, b->feat_cap * sizeof(float), w, rows_min, cudaMemcpyDeviceToDevice));
        FVAD_CUDA_TRY(cudaMemcpy2D(nr, cap * sizeof(float), b->d_feat_ratio, b->feat_cap * sizeof(float), w, b->S, cudaMemcpyDeviceToDevice));
        FVAD_CUDA_TRY(cudaMemcpy2D(nv, cap * sizeof(float), b->d_feat_vad, b->feat_cap * sizeof(float), w, b->S, cudaMemcpyDeviceToDevice));
    }
    cudaFree(b->d_feat_min); cudaFree(b->d_feat_ratio); cudaFree(b->d_feat_vad);
    b->d_feat_min = nm; b->d_feat_ratio = nr; b->d_feat_vad = nv; b->feat_cap = cap;
    return FVAD_OK;
}

static int run_machines(fvad_batch *b, int par, bool with_vad, unsigned long long first_index, size_t J, cudaStream_t st) {
    VadmIo io{};
    io.band = band_buf(b, par); io.band_rows = b->rows; io.band_frame_stride = b->max_stft;
    io.vol_ratio = ratio_buf(b, par); io.vad = with_vad ? vad_buf(b, par) : nullptr; io.ratio_stride = b->max_stft;
    io.n_channels = (int)b->C;
    io.first_index = first_index; io.index_step = (unsigned long long)b->fft_size; io.n_frames = J;
    io.segs = b->d_segs; io.seg_count = b->d_seg_count; io.seg_cap = b->seg_cap;
    io.events = b->d_events; io.event_count = b->d_event_count; io.event_cap = b->event_cap;
    io.trace = (b->trace_flags & FVAD_TRACE_RECORDS) ? b->d_trace : nullptr;
    io.trace_stride = b->max_stft; io.trace_offset = 0;
    if (b->trace_flags & FVAD_KEEP_FEATURES) {
        FVAD_TRY(ensure_feature_capacity(b, b->stft_done + J));
        io.feat_min = b->d_feat_min; io.feat_ratio = b->d_feat_ratio; io.feat_vad = b->d_feat_vad;
        io.feat_stride = b->feat_cap; io.feat_offset = b->stft_done; io.feat_band_stride = b->S * b->feat_cap;
        io.feat_bands = b->bands.n_bands;
    }
    FVAD_TRY(vadm_launch(b->d_params, b->d_states, b->d_rings, b->S, b->n_machines, io, st));
    // the arena counters as of this tile, for the host to look at without waiting (launch_tile)
    FVAD_CUDA_TRY(cudaMemcpyAsync(&b->h_arena_cnt[0], b->d_seg_count, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    FVAD_CUDA_TRY(cudaMemcpyAsync(&b->h_arena_cnt[1], b->d_event_count, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    return FVAD_OK;
}

// One tile: rows of `avail` valid samples at in[row * in_stride] (a work buffer, or the caller's device rows).  Whole
// frames are consumed; the unconsumed tail (< one frame) is copied to the front of `next` (may be null when avail is a
// whole number of frames).
static int process_tile(fvad_batch *b, const float *in, size_t in_stride, size_t avail, float *next, size_t *tail_out) {
    const size_t F = avail / b->frame_len, used = F * b->frame_len;
    size_t J = 0, den_before = b->den_carry;
    cudaStream_t sA = b->use_denoiser ? b->sA : b->stream, sD = b->stream;   // the input side lives on sA
    if (F > 0 && !b->cfg.skip_processing) {                                   // AudioPipeline.zig:153-157
        float *spec = (b->trace_flags & FVAD_TRACE_SPECTRA) ? b->d_spectra : nullptr;
        if (!b->use_denoiser) {
            J = F;
            FVAD_TRY(rms_ratio_launch(in, in_stride, b->S, (int)b->C, F, b->fft_size, b->d_ratio_stft, b->max_stft, 0, sD));
            FVAD_TRY(batch_stft(b, in, in_stride, F, b->d_band, spec, sD));
            FVAD_TRY(run_machines(b, 0, false, b->read_count, J, sD));
        } else {
            // ---- wavefront: stage X of this tile follows stage X of the previous tile (stream order) and
            // stage X-1 of this tile (event); a slot is reused only after its previous tile left stage D.
            const int slot = (int)(b->tile_no % (unsigned)b->n_slots);
            b->tile_no++;
            // stage A (sA): channel RMS ratio + biquad; the tile's PCM is in place (sA has waited for the copy)
            FVAD_CUDA_TRY(cudaStreamWaitEvent(sA, b->evD[slot], 0));
            FVAD_TRY(rms_ratio_launch(in, in_stride, b->S, (int)b->C, F, kRnFrame, b->d_ratio_slot[slot], b->tile_frames, 0, sA));
            FVAD_TRY(rnnoise_stage_a(b->rn, slot, in, in_stride, (int)b->C, F, sA));
            FVAD_CUDA_TRY(cudaEventRecord(b->evA[slot], sA));
            // stage B1 (sB): analysis + pitch correlations, frame-parallel
            FVAD_CUDA_TRY(cudaStreamWaitEvent(b->sB, b->evA[slot], 0));
            FVAD_TRY(rnnoise_stage_b1(b->rn, slot, F * b->C, b->sB));
            FVAD_CUDA_TRY(cudaEventRecord(b->evB1[slot], b->sB));
            // stage B2 (sB2): pitch decision scan, pitch features, cepstral features
            FVAD_CUDA_TRY(cudaStreamWaitEvent(b->sB2, b->evB1[slot], 0));
            FVAD_TRY(rnnoise_stage_b2(b->rn, slot, F * b->C, b->sB2));
            FVAD_CUDA_TRY(cudaEventRecord(b->evB[slot], b->sB2));
            // stage C (sC)
            FVAD_CUDA_TRY(cudaStreamWaitEvent(b->sC, b->evB[slot], 0));
            FVAD_TRY(rnnoise_stage_c(b->rn, slot, F * b->C, b->sC));
            FVAD_CUDA_TRY(cudaEventRecord(b->evC[slot], b->sC));
            // stage D (main stream)
            FVAD_CUDA_TRY(cudaStreamWaitEvent(sD, b->evC[slot], 0));
            FVAD_TRY(copy_rows_launch(b->d_ratio480, b->stride480, fvad_batch::kHist480, b->d_ratio_slot[slot], b->tile_frames, 0,
                                      b->S, F, sD));
            FVAD_TRY(rnnoise_stage_d(b->rn, slot, (int)b->C, F, b->d_den, b->den_cap, b->den_carry, b->d_vad480, nullptr,
                                     b->stride480, fvad_batch::kHist480, sD));
            const size_t den_avail = b->den_carry + F * kRnFrame;
            J = den_avail / b->fft_size;
            if (J > 0) {
                // stage E (sE): the machines read this parity's STFT-rate buffers while stage D of the next tile fills the other
                const int par = (int)(b->stft_tile_no++ & 1);
                FVAD_CUDA_TRY(cudaStreamWaitEvent(sD, b->evE[par], 0));        // the machines of two tiles ago are done with them
                FVAD_TRY(reframe_launch(b->d_ratio480, b->d_vad480, b->stride480, b->base480, fvad_batch::kHist480, b->S,
                                        b->stft_done, J, b->fft_size, ratio_buf(b, par), vad_buf(b, par), b->max_stft, sD));
                FVAD_TRY(batch_stft(b, b->d_den, b->den_cap, J, band_buf(b, par), spec, sD));
                FVAD_CUDA_TRY(cudaEventRecord(b->evS[par], sD));
                FVAD_CUDA_TRY(cudaStreamWaitEvent(b->sE, b->evS[par], 0));
                FVAD_TRY(run_machines(b, par, true, b->stft_done * (unsigned long long)b->fft_size, J, b->sE));
                FVAD_CUDA_TRY(cudaEventRecord(b->evE[par], b->sE));
                b->e_pending = par;
            }
            if (b->trace_flags & kHostTraces) FVAD_TRY(trace_tile(b, F, J, den_before));
            const size_t tail = den_avail - J * b->fft_size;
            if (J > 0 && tail > 0) {
                FVAD_TRY(copy_rows_launch(b->d_den_carry, b->fft_size, 0, b->d_den, b->den_cap, J * b->fft_size, b->rows, tail, sD));
                FVAD_TRY(copy_rows_launch(b->d_den, b->den_cap, 0, b->d_den_carry, b->fft_size, 0, b->rows, tail, sD));
            }
            b->den_carry = tail;
            FVAD_TRY(shift_hist_launch(b->d_ratio480, b->stride480, fvad_batch::kHist480, F, b->S, sD));
            FVAD_TRY(shift_hist_launch(b->d_vad480, b->stride480, fvad_batch::kHist480, F, b->S, sD));
            b->base480 += (long long)F;
            FVAD_CUDA_TRY(cudaEventRecord(b->evD[slot], sD));
        }
        if (!b->use_denoiser && (b->trace_flags & kHostTraces)) FVAD_TRY(trace_tile(b, F, J, 0));
        b->stft_done += J;
        b->dirty = true;
    }
    b->read_count += b->cfg.skip_processing ? 0 : used;
    const size_t consumed = b->cfg.skip_processing ? avail : used;   // skip_processing: samples are only buffered by the reference
    const size_t tail = avail - consumed;
    if (tail > 0) {
        if (!next) return set_error(FVAD_EINVALID_ARGUMENT, "internal: a tile with a partial frame needs a carry target");
        FVAD_TRY(copy_rows_launch(next, b->work_cap, 0, in, in_stride, consumed, b->rows, tail, sA));
    }
    *tail_out = tail;
    return FVAD_OK;
}


// The main stream waits for the machines of the last tile (stage E): whoever synchronises with, or records an event on,
// the batch's stream after a push sees every segment that push produced.
static int join_machines(fvad_batch *b) {
    if (b->e_pending >= 0 && b->sE != b->stream) FVAD_CUDA_TRY(cudaStreamWaitEvent(b->stream, b->evE[b->e_pending], 0));
    b->e_pending = -1;
    return FVAD_OK;
}

static int ensure_trace_buffers(fvad_batch *b) {
    if ((b->trace_flags & FVAD_TRACE_RECORDS) && !b->d_trace) FVAD_TRY(dmalloc(&b->d_trace, b->S * b->max_stft));
    if ((b->trace_flags & FVAD_TRACE_SPECTRA) && !b->d_spectra) FVAD_TRY(dmalloc(&b->d_spectra, b->rows * b->max_stft * ((size_t)b->fft_size / 2 + 1)));
    return FVAD_OK;
}

static int drain_results(fvad_batch *b);
static int extract_clips(fvad_batch *b);

// An arena half full (as of the last tile the device has finished): drain it to the host lists before adding more.
static int arena_check(fvad_batch *b) {
    if (b->h_arena_cnt[0] > b->seg_cap / 2 || b->h_arena_cnt[1] > b->event_cap / 2) return drain_results(b);
    return FVAD_OK;
}

// Samples per row a work buffer takes before its tile is launched.
static size_t tile_samples(const fvad_batch *b) { return b->tile_frames * (size_t)b->frame_len; }

// The filling buffer becomes a tile (stages A..E are enqueued); the other buffer becomes the filling one and starts
// with the partial frame left over, if any.
static int launch_tile(fvad_batch *b) {
    const int cur = b->work_par;
    if (!b->cfg.skip_processing && b->fill_len / b->frame_len == 0) return FVAD_OK;     // no complete frame yet
    cudaStream_t sIn = b->use_denoiser ? b->sA : b->stream;
    FVAD_TRY(arena_check(b));
    if (b->h2d_dirty) {
        FVAD_CUDA_TRY(cudaEventRecord(b->evH[cur], b->sH));
        FVAD_CUDA_TRY(cudaStreamWaitEvent(sIn, b->evH[cur], 0));
        b->h2d_dirty = false;
    }
    size_t tail = 0;
    FVAD_TRY(process_tile(b, b->d_work_buf[cur], b->work_cap, b->fill_len, b->d_work_buf[cur ^ 1], &tail));
    FVAD_CUDA_TRY(cudaEventRecord(b->evW[cur], sIn));                                   // this buffer may be refilled after that
    b->work_par = cur ^ 1;
    b->fill_len = tail;
    b->fill_waited = false;
    return join_machines(b);
}

// Recorder: the m samples per row just appended (device rows at src, row r at src + r * src_stride) also go into the PCM
// history, on the copy stream behind the copy that brought them.
static int history_append(fvad_batch *b, const float *src, size_t src_stride, size_t m) {
    if (!b->rec_on || !m) return FVAD_OK;
    if (!b->hist_fixed && b->hist_len + m > b->hist_cap) {       // everything is kept: the buffer doubles (stop-the-world)
        size_t cap = b->hist_cap ? b->hist_cap : (size_t)1 << 20;
        while (cap < b->hist_len + m) cap *= 2;
        for (cudaStream_t st : {b->sH, b->sA, b->sB, b->sB2, b->sC, b->sE, b->stream}) if (st) FVAD_CUDA_TRY(cudaStreamSynchronize(st));
        float *nh = nullptr;
        FVAD_TRY(dmalloc(&nh, b->rows * cap));
        if (b->hist_len)
            FVAD_CUDA_TRY(cudaMemcpy2D(nh, cap * sizeof(float), b->d_hist, b->hist_cap * sizeof(float), (size_t)b->hist_len * sizeof(float),
                                       b->rows, cudaMemcpyDeviceToDevice));
        cudaFree(b->d_hist);
        b->d_hist = nh; b->hist_cap = cap;
    }
    size_t done = 0;
    while (done < m) {                                           // at most two pieces when the ring wraps
        const size_t pos = (size_t)((b->hist_len + done) % b->hist_cap);
        const size_t n = std::min(m - done, b->hist_cap - pos);
        FVAD_CUDA_TRY(cudaMemcpy2DAsync(b->d_hist + pos, b->hist_cap * sizeof(float), src + done, src_stride * sizeof(float),
                                        n * sizeof(float), b->rows, cudaMemcpyDeviceToDevice, b->sH));
        done += n;
    }
    b->hist_len += m;
    return FVAD_OK;
}

static int ensure_staging(fvad_batch *b) {
    if (b->copy_pool) return FVAD_OK;
    // ~32 MB per slot, whole samples per row, at least 4096 per row
    size_t per_row = ((size_t)32 << 20) / sizeof(float) / b->rows;
    per_row = std::max<size_t>(per_row, 4096);
    per_row = std::min(per_row, tile_samples(b));
    for (int i = 0; i < fvad_batch::kStageSlots; i++) {
        if (cudaHostAlloc((void **)&b->h_stage[i], b->rows * per_row * sizeof(float), cudaHostAllocDefault) != cudaSuccess)
            return set_error(FVAD_EOUT_OF_MEMORY, "pinned staging slot of %zu bytes: %s", b->rows * per_row * sizeof(float),
                             cudaGetErrorString(cudaGetLastError()));
        FVAD_CUDA_TRY(cudaEventCreateWithFlags(&b->evStage[i], cudaEventDisableTiming));
    }
    b->stage_samples = per_row;
    unsigned hw = std::thread::hardware_concurrency();
    int nt = (int)std::min<unsigned>(hw ? hw : 4, 8);
    if (const char *e = debug_knob("copy_threads")) nt = std::max(1, atoi(e));
    b->copy_pool = new (std::nothrow) HostCopyPool(nt);
    if (!b->copy_pool) return set_error(FVAD_EOUT_OF_MEMORY, "host allocation failed");
    return FVAD_OK;
}

// Before the first copy into the filling buffer since it changed hands: its previous tile must have been read.
static int fill_begin(fvad_batch *b) {
    if (!b->fill_waited) {
        FVAD_CUDA_TRY(cudaStreamWaitEvent(b->sH, b->evW[b->work_par], 0));
        b->fill_waited = true;
    }
    return FVAD_OK;
}

extern "C" int fvad_batch_push(fvad_batch *b, const float *const *pcm, size_t n, uint64_t *first_index) {
    if (!b || (!pcm && n)) return set_error(FVAD_EINVALID_ARGUMENT, "fvad_batch_push: null argument");
    DeviceGuard g(b->device);
    if (!g.ok) return set_error(FVAD_ECUDA, "cannot select device %d", b->device);
    if (first_index) *first_index = b->total_write_count;
    FVAD_TRY(ensure_trace_buffers(b));
    // equally strided rows (every caller that keeps its streams in one array): one 2-D copy per tile instead of one
    // copy per row.  cudaMemcpy2DAsync limits the pitch to the device's memPitch (2 GiB - 1).
    ptrdiff_t stride = 0;
    bool strided = n > 0;
    if (strided && b->rows > 1) {
        stride = pcm[1] - pcm[0];
        for (size_t r = 2; r < b->rows && strided; r++) strided = (pcm[r] - pcm[r - 1]) == stride;
        strided = strided && stride >= (ptrdiff_t)n && (size_t)stride * sizeof(float) < ((size_t)1 << 31);
    } else if (strided) stride = (ptrdiff_t)n;
    // pageable rows go through the pinned staging slots (the copy threads move them there); pinned / registered rows are
    // copied where they lie
    bool pageable = false;
    if (n) {
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, pcm[0]) != cudaSuccess) { cudaGetLastError(); pageable = true; }
        else pageable = at.type == cudaMemoryTypeUnregistered;
    }
    if (pageable) FVAD_TRY(ensure_staging(b));
    for (size_t off = 0; off < n;) {
        const size_t m = std::min(tile_samples(b) - b->fill_len, n - off);
        FVAD_TRY(fill_begin(b));
        float *dst = b->d_work_buf[b->work_par] + b->fill_len;
        if (pageable) {
            for (size_t sub = 0; sub < m; sub += b->stage_samples) {
                const size_t len = std::min(b->stage_samples, m - sub);
                const int slot = b->stage_next;
                b->stage_next = (slot + 1) % fvad_batch::kStageSlots;
                FVAD_CUDA_TRY(cudaEventSynchronize(b->evStage[slot]));          // the slot's previous contents are on the device
                float *stage = b->h_stage[slot];
                const size_t src_off = off + sub, ss = b->stage_samples;
                // split long rows so that few-row batches still use every thread
                const size_t parts = b->rows >= 16 ? 1 : 8;
                b->copy_pool->parallel_for(b->rows * parts, [&](size_t i) {
                    const size_t r = i / parts, q = i - r * parts;
                    const size_t a = len * q / parts, z = len * (q + 1) / parts;
                    memcpy(stage + r * ss + a, pcm[r] + src_off + a, (z - a) * sizeof(float));
                });
                FVAD_CUDA_TRY(cudaMemcpy2DAsync(dst + sub, b->work_cap * sizeof(float), stage, ss * sizeof(float), len * sizeof(float),
                                                b->rows, cudaMemcpyHostToDevice, b->sH));
                FVAD_CUDA_TRY(cudaEventRecord(b->evStage[slot], b->sH));
            }
        } else if (strided) {
            FVAD_CUDA_TRY(cudaMemcpy2DAsync(dst, b->work_cap * sizeof(float), pcm[0] + off, (size_t)stride * sizeof(float),
                                            m * sizeof(float), b->rows, cudaMemcpyHostToDevice, b->sH));
        } else {
            for (size_t r = 0; r < b->rows; r++)
                FVAD_CUDA_TRY(cudaMemcpyAsync(dst + r * b->work_cap, pcm[r] + off, sizeof(float) * m, cudaMemcpyHostToDevice, b->sH));
        }
        FVAD_TRY(history_append(b, dst, b->work_cap, m));
        b->h2d_dirty = true;
        b->fill_len += m;
        off += m;
        if (b->fill_len == tile_samples(b)) FVAD_TRY(launch_tile(b));
    }
    b->total_write_count += n;
    // input is borrowed only until return (AudioPipeline.pushSamples copies into its ring): wait for the copies
    // only; the stages keep running behind the caller's back until fvad_batch_sync.  Pageable input has already left the
    // caller's memory (it sits in the staging slots).
    if (!pageable) FVAD_CUDA_TRY(cudaStreamSynchronize(b->sH));
    return FVAD_OK;
}

extern "C" int fvad_batch_push_device(fvad_batch *b, const float *d_pcm, size_t stream_stride, size_t channel_stride,
                                      size_t n, uint64_t *first_index) {
    if (!b || (!d_pcm && n)) return set_error(FVAD_EINVALID_ARGUMENT, "fvad_batch_push_device: null argument");
    DeviceGuard g(b->device);
    if (!g.ok) return set_error(FVAD_ECUDA, "cannot select device %d", b->device);
    if (first_index) *first_index = b->total_write_count;
    FVAD_TRY(ensure_trace_buffers(b));
    // rows (stream, channel) at one stride: a whole tile is read where it lies, no staging copy
    const bool uniform = b->C == 1 || stream_stride == b->C * channel_stride;
    const size_t row_stride = b->C == 1 ? stream_stride : channel_stride;
    const bool pitch_ok = stream_stride * sizeof(float) < ((size_t)1 << 31);
    for (size_t off = 0; off < n;) {
        if (b->fill_len == 0 && uniform && n - off >= tile_samples(b)) {
            size_t tail = 0;
            FVAD_TRY(arena_check(b));
            FVAD_TRY(history_append(b, d_pcm + off, row_stride, tile_samples(b)));
            FVAD_TRY(process_tile(b, d_pcm + off, row_stride, tile_samples(b), nullptr, &tail));
            FVAD_TRY(join_machines(b));
            off += tile_samples(b);
            continue;
        }
        const size_t m = std::min(tile_samples(b) - b->fill_len, n - off);
        FVAD_TRY(fill_begin(b));
        float *dst = b->d_work_buf[b->work_par] + b->fill_len;
        if (pitch_ok) {
            for (size_t c = 0; c < b->C; c++)          // channel c of every stream: one 2-D device-to-device copy
                FVAD_CUDA_TRY(cudaMemcpy2DAsync(dst + c * b->work_cap, b->C * b->work_cap * sizeof(float),
                                                d_pcm + c * channel_stride + off, stream_stride * sizeof(float), m * sizeof(float),
                                                b->S, cudaMemcpyDeviceToDevice, b->sH));
        } else {
            FVAD_TRY(gather_rows_launch(b->d_work_buf[b->work_par], b->work_cap, b->fill_len, d_pcm, stream_stride, channel_stride,
                                        off, b->S, (int)b->C, m, b->sH));
        }
        FVAD_TRY(history_append(b, dst, b->work_cap, m));
        b->h2d_dirty = true;
        b->fill_len += m;
        off += m;
        if (b->fill_len == tile_samples(b)) FVAD_TRY(launch_tile(b));
    }
    b->total_write_count += n;
    return FVAD_OK;
}

static int collect(fvad_batch *b) {
    FVAD_TRY(launch_tile(b));                                  // whatever complete frames the filling buffer holds
    FVAD_TRY(drain_results(b));
    return b->rec_on ? extract_clips(b) : FVAD_OK;             // (drain_results returns early when nothing new was decided)
}

// audio_utils.rmsVolume (src/audio_utils.zig:14-24): f32 sum of squares in order, mean, sqrt
static float rms_volume_host(const float *x, size_t n) {
    float sum = 0.0f;
    for (size_t i = 0; i < n; i++) sum += x[i] * x[i];
    return sqrtf(sum / (float)n);
}

// Recorder: walk the main machine's new events (beginCapture / endCapture, VAD.zig:362-373), then cut every clip whose end
// the stream has reached out of the PCM history (Recorder.finalize -> segmentToAudioBuffer, Recorder.zig:87-146).
static int extract_clips(fvad_batch *b) {
    if (!b->rec_on) return FVAD_OK;
    for (size_t s = 0; s < b->S; s++) {
        fvad_batch::RecState &rs = b->rec_state[s];
        const auto &ev = b->events[s];
        for (; rs.next_event < ev.size(); rs.next_event++) {
            const fvad_capture_event &e = ev[rs.next_event];
            if (e.recording_state == FVAD_REC_STARTED) { rs.active = true; rs.start = e.sample_number; }        // Recorder.start
            else if (e.recording_state == FVAD_REC_ABORTED) rs.active = false;                                   // finalize(0, false)
            else if (e.recording_state == FVAD_REC_COMPLETED && rs.active) {
                rs.active = false;
                if (e.sample_number > rs.start) b->rec_pending.push_back({s, rs.start, e.sample_number});
            }
        }
    }
    // ready = the stream has been pushed up to the clip's end (AudioPipeline.maybeRecordBuffer); keep the others waiting
    std::vector<fvad_batch::PendingClip> ready, wait;
    for (auto &c : b->rec_pending) (c.end <= b->total_write_count ? ready : wait).push_back(c);
    b->rec_pending.swap(wait);
    if (ready.empty()) return FVAD_OK;
    std::stable_sort(ready.begin(), ready.end(), [](const fvad_batch::PendingClip &a, const fvad_batch::PendingClip &c) {
        return a.stream != c.stream ? a.stream < c.stream : a.end < c.end; });
    FVAD_CUDA_TRY(cudaStreamSynchronize(b->sH));                 // the history holds everything pushed
    const size_t C = b->C;
    const size_t stage_max = (size_t)1 << 26;                    // floats per gather round (256 MB)
    size_t i0 = 0;
    while (i0 < ready.size()) {
        // a round: as many clips as fit the staging buffer (a single clip larger than that gets a round of its own)
        std::vector<ClipPiece> pieces;
        size_t i1 = i0, total = 0, max_len = 0;
        while (i1 < ready.size()) {
            const auto &c = ready[i1];
            const size_t len = (size_t)(c.end - c.start);
            const bool lost = b->hist_fixed && c.start + b->hist_cap < b->hist_len;
            const size_t need = lost ? 0 : len * C;
            if (i1 > i0 && total + need > stage_max) break;
            if (!lost) {
                for (size_t ch = 0; ch < C; ch++) pieces.push_back({c.stream * C + ch, c.start, len, total + ch * len});
                max_len = std::max(max_len, len);
            }
            total += need;
            i1++;
        }
        std::vector<float> host(total);
        if (total) {
            if (pieces.size() > b->clip_pieces_cap) {
                cudaFree(b->d_clip_pieces); b->d_clip_pieces = nullptr; b->clip_pieces_cap = 0;
                FVAD_TRY(dmalloc(&b->d_clip_pieces, pieces.size() * 2));
                b->clip_pieces_cap = pieces.size() * 2;
            }
            if (total > b->clip_stage_cap) {
                cudaFree(b->d_clip_stage); b->d_clip_stage = nullptr; b->clip_stage_cap = 0;
                FVAD_TRY(dmalloc(&b->d_clip_stage, total));
                b->clip_stage_cap = total;
            }
            FVAD_CUDA_TRY(cudaMemcpyAsync(b->d_clip_pieces, pieces.data(), sizeof(ClipPiece) * pieces.size(), cudaMemcpyHostToDevice, b->stream));
            FVAD_TRY(gather_clips_launch(b->d_clip_pieces, pieces.size(), max_len, b->d_hist, b->hist_cap, b->d_clip_stage, b->stream));
            FVAD_CUDA_TRY(cudaMemcpyAsync(host.data(), b->d_clip_stage, sizeof(float) * total, cudaMemcpyDeviceToHost, b->stream));
            FVAD_CUDA_TRY(cudaStreamSynchronize(b->stream));
        }
        size_t off = 0;
        for (size_t i = i0; i < i1; i++) {
            const auto &c = ready[i];
            const size_t len = (size_t)(c.end - c.start);
            const bool lost = b->hist_fixed && c.start + b->hist_cap < b->hist_len;
            fvad_batch::StoredClip sc;
            sc.meta.global_start_frame_number = c.start;
            sc.meta.length = len;
            sc.meta.sample_rate = b->cfg.sample_rate;
            sc.meta.duration_seconds = (float)len / (float)b->cfg.sample_rate;     // Recorder.zig:139
            sc.meta.channel = 0; sc.meta.lost = lost ? 1 : 0; sc.meta.pcm = nullptr;
            if (!lost) {
                size_t best = 0;
                float best_vol = 9999.0f;                                           // Recorder.findBestChannel, Recorder.zig:103-119
                for (size_t ch = 0; ch < C; ch++) {
                    const float vol = rms_volume_host(host.data() + off + ch * len, len);
                    if (vol < best_vol) { best = ch; best_vol = vol; }
                }
                sc.meta.channel = best;
                sc.pcm.assign(host.begin() + off + best * len, host.begin() + off + (best + 1) * len);
                off += len * C;
            }
            if (b->on_recording) {
                fvad_clip view = sc.meta;
                view.pcm = lost ? nullptr : sc.pcm.data();
                b->on_recording(b->on_recording_ctx, c.stream, &view);
            }
            b->rec_clips[c.stream].push_back(std::move(sc));
        }
        i0 = i1;
    }
    return FVAD_OK;
}

static int drain_results(fvad_batch *b) {
    FVAD_TRY(join_machines(b));
    FVAD_CUDA_TRY(cudaStreamSynchronize(b->stream));
    if (!b->dirty) return FVAD_OK;
    unsigned ns = 0, ne = 0;
    FVAD_CUDA_TRY(cudaMemcpy(&ns, b->d_seg_count, sizeof ns, cudaMemcpyDeviceToHost));
    FVAD_CUDA_TRY(cudaMemcpy(&ne, b->d_event_count, sizeof ne, cudaMemcpyDeviceToHost));
    if (ns > b->seg_cap || ne > b->event_cap)
        return set_error(FVAD_EOUT_OF_MEMORY, "segment/event arena overflow (%u/%u segments, %u/%u events) although it is drained "
                         "when half full: more than one result per three STFT frames per machine?", ns, b->seg_cap, ne, b->event_cap);
    if (ns) {
        std::vector<SegRecord> tmp(ns);
        FVAD_CUDA_TRY(cudaMemcpy(tmp.data(), b->d_segs, sizeof(SegRecord) * ns, cudaMemcpyDeviceToHost));
        std::stable_sort(tmp.begin(), tmp.end(), [](const SegRecord &a, const SegRecord &c) {
            if (a.stream != c.stream) return a.stream < c.stream;
            if (a.machine != c.machine) return a.machine < c.machine;
            return a.seg.sample_to < c.seg.sample_to;      // a machine emits in increasing end time
        });
        for (auto &r : tmp) b->segs[(size_t)r.stream * b->n_machines + r.machine].push_back(r.seg);
        FVAD_CUDA_TRY(cudaMemset(b->d_seg_count, 0, sizeof(unsigned)));
    }
    if (ne) {
        std::vector<EventRecord> tmp(ne);
        FVAD_CUDA_TRY(cudaMemcpy(tmp.data(), b->d_events, sizeof(EventRecord) * ne, cudaMemcpyDeviceToHost));
        std::stable_sort(tmp.begin(), tmp.end(), [](const EventRecord &a, const EventRecord &c) {
            if (a.stream != c.stream) return a.stream < c.stream;
            return a.ev.at_index < c.ev.at_index;
        });
        for (auto &r : tmp) b->events[r.stream].push_back(r.ev);
        FVAD_CUDA_TRY(cudaMemset(b->d_event_count, 0, sizeof(unsigned)));
    }
    b->h_arena_cnt[0] = b->h_arena_cnt[1] = 0;
    b->dirty = false;
    return extract_clips(b);
}

extern "C" int fvad_batch_sync(fvad_batch *b) {
    if (!b) return set_error(FVAD_EINVALID_ARGUMENT, "null batch");
    DeviceGuard g(b->device);
    return collect(b);
}

extern "C" int fvad_batch_reset(fvad_batch *b) {
    if (!b) return set_error(FVAD_EINVALID_ARGUMENT, "null batch");
    DeviceGuard g(b->device);
    for (cudaStream_t st : {b->sH, b->sA, b->sB, b->sB2, b->sC, b->sE, b->stream}) FVAD_CUDA_TRY(cudaStreamSynchronize(st));
    FVAD_TRY(upload_machine_state(b));
    if (b->use_denoiser) {
        FVAD_TRY(rnnoise_batch_reset(b->rn, b->stream));
        FVAD_CUDA_TRY(cudaMemsetAsync(b->d_ratio480, 0, sizeof(float) * b->S * b->stride480, b->stream));
        FVAD_CUDA_TRY(cudaMemsetAsync(b->d_vad480, 0, sizeof(float) * b->S * b->stride480, b->stream));
    }
    b->fill_len = 0; b->fill_waited = false; b->h2d_dirty = false;
    b->hist_len = 0;
    if (b->rec_on) {
        b->rec_state.assign(b->S, fvad_batch::RecState());
        b->rec_pending.clear();
        for (auto &v : b->rec_clips) v.clear();
    }
    b->den_carry = 0; b->base480 = 0; b->tile_no = 0; b->stft_tile_no = 0; b->e_pending = -1;
    b->work_par = 0;
    b->total_write_count = b->read_count = b->stft_done = 0;
    for (auto &v : b->segs) v.clear();
    for (auto &v : b->events) v.clear();
    for (auto &v : b->tr_records) v.clear();
    for (auto &v : b->tr_denoised) v.clear();
    for (auto &v : b->tr_vad) v.clear();
    for (auto &v : b->tr_spectra) v.clear();
    b->dirty = false;
    FVAD_CUDA_TRY(cudaStreamSynchronize(b->stream));
    return FVAD_OK;
}

extern "C" int fvad_batch_segments(fvad_batch *b, size_t stream, int k, const fvad_segment **segs, size_t *n) {
    if (!b || !segs || !n || stream >= b->S || k < 0 || k >= b->n_machines)
        return set_error(FVAD_EINVALID_ARGUMENT, "fvad_batch_segments: bad argument");
    FVAD_TRY(fvad_batch_sync(b));
    auto &v = b->segs[stream * b->n_machines + k];
    *segs = v.data(); *n = v.size();
    return FVAD_OK;
}

extern "C" int fvad_batch_events(fvad_batch *b, size_t stream, const fvad_capture_event **ev, size_t *n) {
    if (!b || !ev || !n || stream >= b->S) return set_error(FVAD_EINVALID_ARGUMENT, "fvad_batch_events: bad argument");
    FVAD_TRY(fvad_batch_sync(b));
    *ev = b->events[stream].data(); *n = b->events[stream].size();
    return FVAD_OK;
}

extern "C" int fvad_batch_counts(fvad_batch *b, uint64_t *total_write_count, uint64_t *pipeline_read_count) {
    if (!b) return set_error(FVAD_EINVALID_ARGUMENT, "null batch");
    if (total_write_count) *total_write_count = b->total_write_count;
    // VAD.run consumes every whole frame before pushSamples returns (VAD.zig:214-251); here the frames of the filling
    // buffer are consumed at the next tile launch, which does not change the count the reference reports
    if (pipeline_read_count)
        *pipeline_read_count = b->cfg.skip_processing ? 0 : (b->total_write_count / b->frame_len) * (uint64_t)b->frame_len;
    return FVAD_OK;
}

// ---- alt machines after the fact, on the device (BASELINE configs[4]) ------------------------------------------------
template <class T> static int grow(T **p, size_t *cap, size_t need) {
    if (need <= *cap) return FVAD_OK;
    cudaFree(*p); *p = nullptr; *cap = 0;
    FVAD_TRY(dmalloc(p, need));
    *cap = need;
    return FVAD_OK;
}

extern "C" int fvad_batch_sweep(fvad_batch *b, const fvad_vad_machine_config *configs, size_t n_configs, size_t max_segs,
                                fvad_segment *segs, uint32_t *counts, int outputs_on_device) {
    if (!b || !configs || !n_configs || !max_segs || !segs || !counts)
        return set_error(FVAD_EINVALID_ARGUMENT, "fvad_batch_sweep: null/zero argument");
    if (!(b->trace_flags & FVAD_KEEP_FEATURES))
        return set_error(FVAD_EINVALID_ARGUMENT, "fvad_batch_sweep: the batch does not keep its features (fvad_batch_set_trace with "
                         "FVAD_KEEP_FEATURES before the first push)");
    DeviceGuard g(b->device);
    if (!g.ok) return set_error(FVAD_ECUDA, "cannot select device %d", b->device);
    FVAD_TRY(collect(b));                                       // everything pushed so far has reached the feature planes
    std::vector<VadmParams> params(n_configs);
    for (size_t k = 0; k < n_configs; k++) {
        size_t nd = 0;
        FVAD_TRY(vadm_params_from_config(configs[k], b->cfg.sample_rate, (size_t)b->fft_size, &params[k], &nd));
        int lo = 0, hi = 0;
        FVAD_TRY(freq_to_bin(b->cfg.sample_rate, (size_t)b->fft_size, configs[k].speech_min_freq, &lo));
        FVAD_TRY(freq_to_bin(b->cfg.sample_rate, (size_t)b->fft_size, configs[k].speech_max_freq, &hi));
        int band = -1;
        for (int i = 0; i < b->bands.n_bands; i++) if (b->bands.lo[i] == lo && b->bands.hi[i] == hi) band = i;
        if (band < 0)
            return set_error(FVAD_EUNSUPPORTED, "fvad_batch_sweep: config %zu reads bins %d..%d, a band this batch keeps no sums of "
                             "(create it with a machine config of that band)", k, lo, hi);
        params[k].band = band;
    }
    const size_t n_scans = b->S * n_configs, n_rec = (size_t)b->stft_done;
    int max_long = 0, max_short = 0, max_ratio = 0;
    vadm_sweep_rows(params.data(), n_configs, &max_long, &max_short, &max_ratio);
    FVAD_TRY(grow(&b->d_sw_params, &b->sw_params_cap, n_configs));
    FVAD_TRY(grow(&b->d_sw_rings, &b->sw_rings_cap, (size_t)(max_long + max_short + max_ratio) * n_scans));
    fvad_segment *d_segs = segs; uint32_t *d_counts = counts;
    if (!outputs_on_device) {
        FVAD_TRY(grow(&b->d_sw_segs, &b->sw_segs_cap, n_scans * max_segs));
        FVAD_TRY(grow(&b->d_sw_counts, &b->sw_counts_cap, n_scans));
        d_segs = b->d_sw_segs; d_counts = b->d_sw_counts;
    }
    if (!n_rec) {
        if (outputs_on_device) FVAD_CUDA_TRY(cudaMemsetAsync(counts, 0, sizeof(uint32_t) * n_scans, b->stream));
        else memset(counts, 0, sizeof(uint32_t) * n_scans);
        return FVAD_OK;
    }
    FVAD_CUDA_TRY(cudaMemcpyAsync(b->d_sw_params, params.data(), sizeof(VadmParams) * n_configs, cudaMemcpyHostToDevice, b->stream));
    VadmSweepIo io{};
    io.min_volume = b->d_feat_min; io.band_stride = b->S * b->feat_cap; io.vol_ratio = b->d_feat_ratio;
    io.vad = b->use_denoiser ? b->d_feat_vad : nullptr;        // fft_input.vad is null without the denoiser (VAD.zig:306-312)
    io.rec_stride = b->feat_cap; io.n_records = n_rec;
    io.first_index = 0; io.index_step = (unsigned long long)b->fft_size;
    io.segs = d_segs; io.max_segs = max_segs; io.counts = d_counts; io.fallbacks = nullptr;
    FVAD_TRY(vadm_sweep_launch(b->d_sw_params, b->d_sw_rings, max_long, max_short, b->S, (int)n_configs, io, b->stream));
    if (!outputs_on_device) {
        FVAD_CUDA_TRY(cudaMemcpyAsync(counts, d_counts, sizeof(uint32_t) * n_scans, cudaMemcpyDeviceToHost, b->stream));
        FVAD_CUDA_TRY(cudaMemcpyAsync(segs, d_segs, sizeof(fvad_segment) * n_scans * max_segs, cudaMemcpyDeviceToHost, b->stream));
    }
    FVAD_CUDA_TRY(cudaStreamSynchronize(b->stream));            // `params` is pageable: it must outlive the copy
    return FVAD_OK;
}

extern "C" int fvad_batch_features(fvad_batch *b, size_t stream, fvad_stft_record *out, size_t cap, size_t *n) {
    if (!b || stream >= b->S) return set_error(FVAD_EINVALID_ARGUMENT, "fvad_batch_features: bad argument");
    if (!(b->trace_flags & FVAD_KEEP_FEATURES))
        return set_error(FVAD_EINVALID_ARGUMENT, "fvad_batch_features: the batch does not keep its features (FVAD_KEEP_FEATURES)");
    DeviceGuard g(b->device);
    if (!g.ok) return set_error(FVAD_ECUDA, "cannot select device %d", b->device);
    FVAD_TRY(collect(b));
    const size_t have = (size_t)b->stft_done, m = std::min(cap, have);
    if (n) *n = have;
    if (!out || !m) return FVAD_OK;
    const int band0 = b->h_params[0].band;                      // the MAIN machine's band, as in the records trace
    std::vector<float> mv(m), rt(m), vd(m);
    FVAD_CUDA_TRY(cudaMemcpy(mv.data(), b->d_feat_min + ((size_t)band0 * b->S + stream) * b->feat_cap, sizeof(float) * m, cudaMemcpyDeviceToHost));
    FVAD_CUDA_TRY(cudaMemcpy(rt.data(), b->d_feat_ratio + stream * b->feat_cap, sizeof(float) * m, cudaMemcpyDeviceToHost));
    FVAD_CUDA_TRY(cudaMemcpy(vd.data(), b->d_feat_vad + stream * b->feat_cap, sizeof(float) * m, cudaMemcpyDeviceToHost));
    for (size_t f = 0; f < m; f++) {
        out[f].index = (uint64_t)f * (uint64_t)b->fft_size;
        out[f].min_volume = mv[f]; out[f].vol_ratio = rt[f]; out[f].rnn_vad = vd[f];
        out[f].has_vad = b->use_denoiser ? 1.0f : 0.0f;
    }
    return FVAD_OK;
}

extern "C" int fvad_batch_feature_count(fvad_batch *b, size_t *n_records) {
    if (!b || !n_records) return set_error(FVAD_EINVALID_ARGUMENT, "null argument");
    *n_records = (size_t)b->stft_done + 0;
    return FVAD_OK;
}

extern "C" int fvad_batch_set_recording(fvad_batch *b, size_t history_samples, fvad_on_recording_fn on_recording, void *ctx) {
    if (!b) return set_error(FVAD_EINVALID_ARGUMENT, "null batch");
    if (b->total_write_count) return set_error(FVAD_EINVALID_ARGUMENT, "enable the recorder before the first push");
    DeviceGuard g(b->device);
    if (!g.ok) return set_error(FVAD_ECUDA, "cannot select device %d", b->device);
    cudaFree(b->d_hist); b->d_hist = nullptr; b->hist_cap = 0; b->hist_len = 0;
    b->hist_fixed = history_samples != 0;
    if (b->hist_fixed) {
        FVAD_TRY(dmalloc(&b->d_hist, b->rows * history_samples));
        b->hist_cap = history_samples;
    }
    b->rec_on = true;
    b->on_recording = on_recording; b->on_recording_ctx = ctx;
    b->rec_state.assign(b->S, fvad_batch::RecState());
    b->rec_pending.clear();
    b->rec_clips.assign(b->S, {}); b->rec_view.assign(b->S, {}); b->rec_held.assign(b->S, {});
    return FVAD_OK;
}

extern "C" int fvad_batch_clips(fvad_batch *b, size_t stream, const fvad_clip **clips, size_t *n) {
    if (!b || !clips || !n || stream >= b->S) return set_error(FVAD_EINVALID_ARGUMENT, "fvad_batch_clips: bad argument");
    if (!b->rec_on) return set_error(FVAD_EINVALID_ARGUMENT, "fvad_batch_clips: the recorder is off (fvad_batch_set_recording)");
    FVAD_TRY(fvad_batch_sync(b));
    b->rec_held[stream].swap(b->rec_clips[stream]);            // what was handed out before is released now
    b->rec_clips[stream].clear();
    auto &view = b->rec_view[stream];
    view.clear();
    for (auto &sc : b->rec_held[stream]) {
        fvad_clip c = sc.meta;
        c.pcm = sc.meta.lost ? nullptr : sc.pcm.data();
        view.push_back(c);
    }
    *clips = view.data(); *n = view.size();
    return FVAD_OK;
}

extern "C" void *fvad_batch_cuda_stream(fvad_batch *b) { return b ? (void *)b->stream : nullptr; }

extern "C" int fvad_batch_set_trace(fvad_batch *b, unsigned flags) {
    if (!b) return set_error(FVAD_EINVALID_ARGUMENT, "null batch");
    if (b->total_write_count) return set_error(FVAD_EINVALID_ARGUMENT, "set the trace flags before the first push");
    b->trace_flags = flags;
    return FVAD_OK;
}

template <class T> static int copy_out(const std::vector<T> &v, T *out, size_t cap, size_t *n) {
    if (n) *n = v.size();
    if (out && cap) memcpy(out, v.data(), sizeof(T) * std::min(cap, v.size()));
    return FVAD_OK;
}
extern "C" int fvad_batch_trace_records(fvad_batch *b, size_t s, fvad_stft_record *out, size_t cap, size_t *n) {
    if (!b || s >= b->S) return set_error(FVAD_EINVALID_ARGUMENT, "bad argument");
    FVAD_TRY(fvad_batch_sync(b));
    return copy_out(b->tr_records[s], out, cap, n);
}
extern "C" int fvad_batch_trace_denoised(fvad_batch *b, size_t s, size_t c, float *out, size_t cap, size_t *n) {
    if (!b || s >= b->S || c >= b->C) return set_error(FVAD_EINVALID_ARGUMENT, "bad argument");
    FVAD_TRY(fvad_batch_sync(b));
    return copy_out(b->tr_denoised[s * b->C + c], out, cap, n);
}
extern "C" int fvad_batch_trace_rnn_vad(fvad_batch *b, size_t s, float *out, size_t cap, size_t *n) {
    if (!b || s >= b->S) return set_error(FVAD_EINVALID_ARGUMENT, "bad argument");
    FVAD_TRY(fvad_batch_sync(b));
    return copy_out(b->tr_vad[s], out, cap, n);
}
extern "C" int fvad_batch_trace_spectra(fvad_batch *b, size_t s, size_t c, float *out, size_t cap, size_t *n) {
    if (!b || s >= b->S || c >= b->C) return set_error(FVAD_EINVALID_ARGUMENT, "bad argument");
    FVAD_TRY(fvad_batch_sync(b));
    return copy_out(b->tr_spectra[s * b->C + c], out, cap, n);
}

// ---- single pipeline = batch of 1 (AudioPipeline.init / pushSamples / deinit) ---------------
extern "C" int fvad_pipeline_create(const fvad_config *cfg, fvad_pipeline **out) { return fvad_batch_create(cfg, 1, -1, out); }
extern "C" int fvad_pipeline_push(fvad_pipeline *p, const float *const *channel_pcm, size_t n, uint64_t *first_index) {
    return fvad_batch_push(p, channel_pcm, n, first_index);
}
extern "C" int fvad_pipeline_segments(fvad_pipeline *p, int k, const fvad_segment **segs, size_t *n) {
    return fvad_batch_segments(p, 0, k, segs, n);
}
extern "C" void fvad_pipeline_destroy(fvad_pipeline *p) { fvad_batch_destroy(p); }
