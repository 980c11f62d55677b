// runtime.cu — error text, launch counter, version (no kernels).
#include "common.cuh"
#include <atomic>
#include <cstdarg>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace lodgs {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

// ---- optional per-kernel timing: an event is recorded after every launch (and at every API entry); the time
// between consecutive events on the stream is attributed to the kernel that was just launched.
struct ProfMark { cudaEvent_t ev; const char* name; };
static bool g_prof_on = false;
static std::vector<ProfMark> g_marks;
static std::vector<cudaEvent_t> g_pool;
static thread_local cudaStream_t g_prof_stream = 0;
static std::mutex g_prof_mu;

static void prof_record(const char* name) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    cudaEvent_t ev;
    if (!g_pool.empty()) { ev = g_pool.back(); g_pool.pop_back(); }
    else if (cudaEventCreate(&ev) != cudaSuccess) return;
    cudaEventRecord(ev, g_prof_stream);
    g_marks.push_back({ev, name});
}
void prof_stage_begin(cudaStream_t s) {
    g_prof_stream = s;
    if (g_prof_on) prof_record(nullptr);
}
void prof_mark(const char* name) {
    if (g_prof_on) prof_record(name);
}
int check_cuda(cudaError_t e, const char* what) {
    set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
    return LODGS_E_CUDA;
}

// ---- mapped pinned result words (see common.cuh)
constexpr int RESULT_WORDS = 8;
static thread_local int32_t* t_words = nullptr;
static thread_local int32_t* t_words_dev = nullptr;
static thread_local cudaEvent_t t_words_evs[MAX_DEVICES] = {};   // an event belongs to the device it was created on
static thread_local cudaEvent_t t_words_ev = nullptr;             // the one recorded by the last publish

__global__ void publish_words_kernel(const int32_t* __restrict__ src, int32_t* __restrict__ host_words, int n) {
    for (int i = 0; i < n; i++) reinterpret_cast<volatile int32_t*>(host_words)[i] = src[i];
    __threadfence_system();
}

int publish_words(const int32_t* src_dev, int n, cudaStream_t s) {
    if (n < 1 || n > RESULT_WORDS) { set_error("publish_words: bad word count %d", n); return LODGS_E_ARG; }
    if (!t_words) LODGS_CUDA(cudaHostAlloc((void**)&t_words, RESULT_WORDS * sizeof(int32_t), cudaHostAllocMapped | cudaHostAllocPortable));
    const int dslot = device_slot();
    if (!t_words_evs[dslot]) LODGS_CUDA(cudaEventCreateWithFlags(&t_words_evs[dslot], cudaEventDisableTiming));
    t_words_ev = t_words_evs[dslot];
    LODGS_CUDA(cudaHostGetDevicePointer((void**)&t_words_dev, t_words, 0));  // per current device
    publish_words_kernel<<<1, 1, 0, s>>>(src_dev, t_words_dev, n);
    LODGS_LAUNCHED("publish_words_kernel");
    LODGS_CUDA(cudaEventRecord(t_words_ev, s));
    return LODGS_OK;
}

const int32_t* wait_words() {
    if (!t_words_ev || cudaEventSynchronize(t_words_ev) != cudaSuccess) return nullptr;
    return t_words;
}

}  // namespace lodgs

extern "C" const char* lodgs_last_error(void) { return lodgs::g_err; }
extern "C" int lodgs_version(void) { return 100; }
extern "C" int lodgs_compiled_sm(void) { return 100; }
extern "C" uint64_t lodgs_launch_count(void) { return lodgs::g_launches.load(); }

extern "C" int lodgs_profile(int enable) {
    std::lock_guard<std::mutex> lk(lodgs::g_prof_mu);
    for (auto& m : lodgs::g_marks) lodgs::g_pool.push_back(m.ev);
    lodgs::g_marks.clear();
    lodgs::g_prof_on = enable != 0;
    return LODGS_OK;
}

extern "C" int lodgs_profile_dump(char* buf, size_t nbytes) {
    if (!buf || nbytes == 0) { lodgs::set_error("lodgs_profile_dump: bad argument"); return LODGS_E_ARG; }
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return lodgs::check_cuda(e, "cudaDeviceSynchronize");
    std::lock_guard<std::mutex> lk(lodgs::g_prof_mu);
    std::map<std::string, std::pair<double, long>> agg;
    std::vector<std::string> order;
    for (size_t i = 1; i < lodgs::g_marks.size(); i++) {
        if (!lodgs::g_marks[i].name) continue;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, lodgs::g_marks[i - 1].ev, lodgs::g_marks[i].ev) != cudaSuccess) { cudaGetLastError(); continue; }
        auto it = agg.find(lodgs::g_marks[i].name);
        if (it == agg.end()) { order.push_back(lodgs::g_marks[i].name); it = agg.emplace(lodgs::g_marks[i].name, std::make_pair(0.0, 0L)).first; }
        it->second.first += ms;
        it->second.second += 1;
    }
    size_t off = 0;
    buf[0] = 0;
    for (auto& n : order) {
        int w = snprintf(buf + off, nbytes - off, "%s %.6f %ld\n", n.c_str(), agg[n].first, agg[n].second);
        if (w < 0 || (size_t)w >= nbytes - off) break;
        off += (size_t)w;
    }
    return LODGS_OK;
}
