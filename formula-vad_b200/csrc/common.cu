// common.cu -- error reporting, launch accounting, misc C-ABI entry points.
#include "common.cuh"
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace fvad {

static thread_local char t_err[1024] = "";
std::atomic<uint64_t> g_launch_count{0};

int set_error(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof t_err, fmt, ap);
    va_end(ap);
    return code;
}

std::atomic<int> g_profile_on{0};
namespace {
struct ProfRec { std::string name; cudaEvent_t a, b; };
std::mutex g_prof_mu;
std::vector<ProfRec> g_prof;
}
void profile_begin(const char *name, cudaStream_t stream) {
    ProfRec r; r.name = name;
    if (cudaEventCreate(&r.a) != cudaSuccess || cudaEventCreate(&r.b) != cudaSuccess) return;
    cudaEventRecord(r.a, stream);
    std::lock_guard<std::mutex> lk(g_prof_mu);
    g_prof.push_back(r);
}
void profile_end(cudaStream_t stream) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    if (!g_prof.empty()) cudaEventRecord(g_prof.back().b, stream);
}

namespace {
std::mutex g_knob_mu;
std::map<std::string, std::string> g_knobs;
}
const char *debug_knob(const char *key) {
    static thread_local std::string copy;
    std::lock_guard<std::mutex> lk(g_knob_mu);
    auto it = g_knobs.find(key);
    if (it == g_knobs.end()) return nullptr;
    copy = it->second;
    return copy.c_str();
}

bool latency_regime(size_t n_chains, int device) {
    if (const char *e = debug_knob("latency_mode")) return atoi(e) != 0;
    // Round 1 (one chain per reserved SM), x real-time latency / throughput rule: 40 chains 28.0k / 22.0k, 80: 47.5k / 37.7k,
    // 110: 48.9k / 47.7k, 148: 49.8k / 54.5k -- the crossover sat near three quarters of the SM count.  Round 2 packs up to
    // four chains on a reserved SM (rnnoise.cu, form 5), ms per 240 s latency / throughput rule: 128 chains 321.9 / 372.7,
    // 148: 370.7 / 384.4, 160: 385.7 / 398.0 -- the rule now holds up to one chain per SM.
    return n_chains <= (size_t)device_sm_count(device);
}

int device_sm_count(int device) {
    static int cache[64];
    if (device >= 0 && device < 64 && cache[device]) return cache[device];
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0) n = 148;
    if (device >= 0 && device < 64) cache[device] = n;
    return n;
}

}  // namespace fvad

extern "C" {

const char *fvad_last_error(void) { return fvad::t_err; }

const char *fvad_status_name(int s) {
    switch (s) {
    case FVAD_OK: return "Ok";
    case FVAD_EINVALID_SAMPLE_RATE: return "InvalidSampleRate";
    case FVAD_EINVALID_FFT_SIZE: return "InvalidFFTSize";
    case FVAD_EINVALID_SAMPLES_LEN: return "InvalidSamplesLength";
    case FVAD_EINVALID_WINDOW_LEN: return "InvalidWindowLength";
    case FVAD_EINVALID_RESULT_LEN: return "InvalidResultLength";
    case FVAD_EOUT_OF_RANGE: return "OutOfRange";
    case FVAD_ENEGATIVE_FREQUENCY: return "NegativeFrequency";
    case FVAD_EINVALID_FRAME_SIZE: return "InvalidFrameSize";
    case FVAD_EINVALID_RESULT_SIZE: return "InvalidResultSize";
    case FVAD_ENOT_INITIALIZED: return "NotInitialized";
    case FVAD_ECHANNEL_COUNT_MISMATCH: return "ChannelCountMismatch";
    case FVAD_EOUT_OF_MEMORY: return "OutOfMemory";
    case FVAD_EINVALID_ARGUMENT: return "InvalidArgument";
    case FVAD_EUNSUPPORTED: return "Unsupported";
    case FVAD_ECUDA: return "CudaError";
    case FVAD_EINSUFFICIENT_SAMPLES: return "InsufficientSamples";
    default: return "Unknown";
    }
}

int fvad_debug_set(const char *key, const char *value) {
    std::lock_guard<std::mutex> lk(fvad::g_knob_mu);
    if (!key) { fvad::g_knobs.clear(); return FVAD_OK; }
    if (!value) fvad::g_knobs.erase(key);
    else fvad::g_knobs[key] = value;
    return FVAD_OK;
}

const char *fvad_version(void) { return "fvad-b200 0.2 (sm_100a)"; }

int fvad_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) return fvad::set_error(FVAD_ECUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
    return n;
}

uint64_t fvad_launch_count(void) { return fvad::g_launch_count.load(); }

void fvad_profile_enable(int on) { fvad::g_profile_on.store(on ? 1 : 0); }

// JSON: {"kernel": {"launches": n, "total_ms": t}, ...} for everything recorded since the last report.
int fvad_profile_report(char *buf, size_t cap) {
    using namespace fvad;
    cudaDeviceSynchronize();
    std::map<std::string, std::pair<long, double>> agg;
    {
        std::lock_guard<std::mutex> lk(g_prof_mu);
        // measurement aid: knob timeline=<file> writes every launch's start / end (ms since the first recorded launch),
        // which shows how the stages of the wavefront overlap
        if (const char *tl = debug_knob("timeline")) {
            if (FILE *f = (buf && cap && !g_prof.empty()) ? fopen(tl, "w") : nullptr) {
                fprintf(f, "kernel,start_ms,end_ms\n");
                for (auto &r : g_prof) {
                    float t0 = 0, t1 = 0;
                    if (cudaEventElapsedTime(&t0, g_prof[0].a, r.a) == cudaSuccess && cudaEventElapsedTime(&t1, g_prof[0].a, r.b) == cudaSuccess)
                        fprintf(f, "%s,%.4f,%.4f\n", r.name.c_str(), t0, t1);
                }
                fclose(f);
            }
        }
        for (auto &r : g_prof) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) { agg[r.name].first++; agg[r.name].second += ms; }
            if (buf && cap) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
        }
        if (buf && cap) g_prof.clear();      // a size query (buf == NULL) leaves the records in place
    }
    std::string out = "{";
    bool first = true;
    for (auto &kv : agg) {
        char tmp[256];
        snprintf(tmp, sizeof tmp, "%s\"%s\": {\"launches\": %ld, \"total_ms\": %.6f}", first ? "" : ", ", kv.first.c_str(), kv.second.first, kv.second.second);
        out += tmp; first = false;
    }
    out += "}";
    if (buf && cap) { snprintf(buf, cap, "%s", out.c_str()); }
    return (int)out.size();
}

void fvad_default_vad_machine_config(fvad_vad_machine_config *c) {   // VADMachine.zig:18-39
    memset(c, 0, sizeof *c);
    c->speech_min_freq = 100;
    c->speech_max_freq = 1500;
    c->long_term_speech_avg_sec = 180;
    c->has_initial_long_term_avg = 1;
    c->initial_long_term_avg = 0.005;
    c->short_term_speech_avg_sec = 0.2f;
    c->speech_threshold_factor = 18;
    c->channel_vol_ratio_avg_sec = 0.5f;
    c->channel_vol_ratio_threshold = 0.5f;
    c->min_consecutive_sec_to_open = 0.2f;
    c->max_speech_gap_sec = 2;
    c->min_vad_duration_sec = 0.7f;
}

void fvad_default_config(fvad_config *c, size_t sample_rate, size_t n_channels) {   // AudioPipeline.zig:20-26, VAD.zig:17-23
    memset(c, 0, sizeof *c);
    c->sample_rate = sample_rate;
    c->n_channels = n_channels;
    c->buffer_length = 0;
    c->vad_config.fft_size = 2048;
    c->vad_config.use_denoiser = 1;
    fvad_default_vad_machine_config(&c->vad_config.vad_machine_config);
    c->vad_config.alt_vad_machine_configs = nullptr;
    c->vad_config.n_alt_vad_machine_configs = 0;
    c->skip_processing = 0;
}

}  // extern "C"
