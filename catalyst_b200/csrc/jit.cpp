// jit.cpp — NVRTC + driver-API plumbing of the specialised tile passes (see jit.hpp).
#include "jit.hpp"

#include <cuda_runtime.h>
#include <dlfcn.h>

#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <thread>

namespace b200 {
namespace jit {

namespace {

// ---- the few driver-API / NVRTC entry points used, resolved with dlopen --------------------------
typedef int CUresult_t;
typedef void *CUmodule_t;
typedef void *CUfunction_t;
typedef void *nvrtcProgram_t;
enum { FUNC_ATTR_NUM_REGS = 4, FUNC_ATTR_MAX_DYNAMIC_SHARED = 8 };

struct Api {
    void *h_cuda = nullptr, *h_nvrtc = nullptr;
    CUresult_t (*cuInit)(unsigned) = nullptr;
    CUresult_t (*cuModuleLoadData)(CUmodule_t *, const void *) = nullptr;
    CUresult_t (*cuModuleGetFunction)(CUfunction_t *, CUmodule_t, const char *) = nullptr;
    CUresult_t (*cuFuncSetAttribute)(CUfunction_t, int, int) = nullptr;
    CUresult_t (*cuFuncGetAttribute)(int *, int, CUfunction_t) = nullptr;
    CUresult_t (*cuOccupancyMaxActiveBlocksPerMultiprocessor)(int *, CUfunction_t, int, size_t) = nullptr;
    CUresult_t (*cuLaunchKernel)(CUfunction_t, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned,
                                 unsigned, void *, void **, void **) = nullptr;
    CUresult_t (*cuGetErrorString)(CUresult_t, const char **) = nullptr;
    int (*nvrtcCreateProgram)(nvrtcProgram_t *, const char *, const char *, int, const char *const *,
                              const char *const *) = nullptr;
    int (*nvrtcCompileProgram)(nvrtcProgram_t, int, const char *const *) = nullptr;
    int (*nvrtcGetCUBINSize)(nvrtcProgram_t, size_t *) = nullptr;
    int (*nvrtcGetCUBIN)(nvrtcProgram_t, char *) = nullptr;
    int (*nvrtcGetProgramLogSize)(nvrtcProgram_t, size_t *) = nullptr;
    int (*nvrtcGetProgramLog)(nvrtcProgram_t, char *) = nullptr;
    int (*nvrtcDestroyProgram)(nvrtcProgram_t *) = nullptr;
    std::string why;
    bool ok = false;
};

template <typename F> bool sym(void *h, const char *name, F &out, std::string &why)
{
    out = reinterpret_cast<F>(dlsym(h, name));
    if (!out) why += std::string(" missing symbol ") + name + ";";
    return out != nullptr;
}

void load_api(Api &a)
{
    const char *cuda_names[] = {"libcuda.so.1", "libcuda.so"};
    for (const char *n : cuda_names)
        if ((a.h_cuda = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!a.h_cuda) {
        a.why = "libcuda.so.1 (driver API) not loadable";
        return;
    }
    const char *env = std::getenv("B200SV_NVRTC");
    const char *nvrtc_names[] = {env ? env : "libnvrtc.so.12", "libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so.12",
                                 "/usr/local/cuda/lib64/libnvrtc.so", "libnvrtc.so"};
    for (const char *n : nvrtc_names)
        if ((a.h_nvrtc = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!a.h_nvrtc) {
        a.why = "libnvrtc.so.12 (NVRTC) not loadable";
        return;
    }
    bool ok = true;
    ok &= sym(a.h_cuda, "cuInit", a.cuInit, a.why);
    ok &= sym(a.h_cuda, "cuModuleLoadData", a.cuModuleLoadData, a.why);
    ok &= sym(a.h_cuda, "cuModuleGetFunction", a.cuModuleGetFunction, a.why);
    ok &= sym(a.h_cuda, "cuFuncSetAttribute", a.cuFuncSetAttribute, a.why);
    ok &= sym(a.h_cuda, "cuFuncGetAttribute", a.cuFuncGetAttribute, a.why);
    ok &= sym(a.h_cuda, "cuOccupancyMaxActiveBlocksPerMultiprocessor", a.cuOccupancyMaxActiveBlocksPerMultiprocessor, a.why);
    ok &= sym(a.h_cuda, "cuLaunchKernel", a.cuLaunchKernel, a.why);
    ok &= sym(a.h_cuda, "cuGetErrorString", a.cuGetErrorString, a.why);
    ok &= sym(a.h_nvrtc, "nvrtcCreateProgram", a.nvrtcCreateProgram, a.why);
    ok &= sym(a.h_nvrtc, "nvrtcCompileProgram", a.nvrtcCompileProgram, a.why);
    ok &= sym(a.h_nvrtc, "nvrtcGetCUBINSize", a.nvrtcGetCUBINSize, a.why);
    ok &= sym(a.h_nvrtc, "nvrtcGetCUBIN", a.nvrtcGetCUBIN, a.why);
    ok &= sym(a.h_nvrtc, "nvrtcGetProgramLogSize", a.nvrtcGetProgramLogSize, a.why);
    ok &= sym(a.h_nvrtc, "nvrtcGetProgramLog", a.nvrtcGetProgramLog, a.why);
    ok &= sym(a.h_nvrtc, "nvrtcDestroyProgram", a.nvrtcDestroyProgram, a.why);
    if (ok && a.cuInit(0) != 0) {
        a.why = "cuInit failed";
        ok = false;
    }
    a.ok = ok;
}

enum State { SEEN = 0, COMPILING = 1, COMPILED = 2, READY = 3, FAILED = 4 };

struct Entry {
    int state = SEEN;
    int sightings = 0;
    Kernel k;
    std::string source;
    std::vector<char> cubin;
};

} // namespace

struct Engine::Impl {
    std::once_flag once;
    Api api;
    std::mutex mu;
    std::condition_variable cv_jobs, cv_done;
    std::map<std::string, std::shared_ptr<Entry>> cache; // key = device + structural key
    std::deque<std::shared_ptr<Entry>> jobs;
    std::vector<std::thread> pool;
    bool stop = false;
    Counters ctr;
    std::map<int, int> sms;

    void compile(Entry &e) // runs without the lock; touches only e.source / e.cubin / e.k
    {
        const auto t0 = std::chrono::steady_clock::now();
        nvrtcProgram_t prog = nullptr;
        std::string err;
        if (api.nvrtcCreateProgram(&prog, e.source.c_str(), "b200sv_pass.cu", 0, nullptr, nullptr) != 0)
            err = "nvrtcCreateProgram failed";
        else {
            const char *opts[] = {"--gpu-architecture=sm_100a", "-lineinfo", "--std=c++17", "-default-device"};
            const int rc = api.nvrtcCompileProgram(prog, 4, opts);
            if (rc != 0) {
                size_t n = 0;
                api.nvrtcGetProgramLogSize(prog, &n);
                std::string log(n, '\0');
                if (n) api.nvrtcGetProgramLog(prog, &log[0]);
                err = "NVRTC: " + log;
            }
            else {
                size_t n = 0;
                api.nvrtcGetCUBINSize(prog, &n);
                e.cubin.resize(n);
                if (n == 0 || api.nvrtcGetCUBIN(prog, e.cubin.data()) != 0) err = "nvrtcGetCUBIN failed";
            }
            api.nvrtcDestroyProgram(&prog);
        }
        e.k.error = err;
        e.k.compile_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }

    void finish_compile(Entry &e) // with the lock
    {
        e.state = e.k.error.empty() ? COMPILED : FAILED;
        ctr.compile_ms += e.k.compile_ms;
        if (e.state == FAILED) {
            ctr.failed++;
            if (std::getenv("B200SV_JIT_VERBOSE")) fprintf(stderr, "[b200sv jit] %s\n", e.k.error.c_str());
        }
    }

    // module load + attributes, on the caller's thread (its CUDA context is current); with the lock
    void make_ready(Entry &e, int cuda_device)
    {
        cudaSetDevice(cuda_device);
        cudaFree(nullptr); // the runtime's primary context becomes current on this thread
        CUmodule_t mod = nullptr;
        CUfunction_t fn = nullptr;
        auto fail = [&](const char *what, CUresult_t rc) {
            const char *s = nullptr;
            api.cuGetErrorString(rc, &s);
            e.k.error = std::string(what) + ": " + (s ? s : "?");
            e.state = FAILED;
            ctr.failed++;
            if (std::getenv("B200SV_JIT_VERBOSE")) fprintf(stderr, "[b200sv jit] %s\n", e.k.error.c_str());
        };
        CUresult_t rc = api.cuModuleLoadData(&mod, e.cubin.data());
        if (rc != 0) return fail("cuModuleLoadData", rc);
        rc = api.cuModuleGetFunction(&fn, mod, e.k.name.c_str());
        if (rc != 0) return fail("cuModuleGetFunction", rc);
        rc = api.cuFuncSetAttribute(fn, FUNC_ATTR_MAX_DYNAMIC_SHARED, e.k.smem_bytes);
        if (rc != 0) return fail("cuFuncSetAttribute(max dynamic shared)", rc);
        int nb = 0, regs = 0;
        api.cuOccupancyMaxActiveBlocksPerMultiprocessor(&nb, fn, e.k.threads, (size_t)e.k.smem_bytes);
        api.cuFuncGetAttribute(&regs, FUNC_ATTR_NUM_REGS, fn);
        e.k.fn = fn;
        e.k.ctas_per_sm = nb > 0 ? nb : 1;
        e.k.num_regs = regs;
        e.cubin.clear();
        e.cubin.shrink_to_fit();
        e.state = READY;
        ctr.compiled++;
        if (std::getenv("B200SV_JIT_VERBOSE"))
            fprintf(stderr, "[b200sv jit] %s: %d regs, %d CTAs/SM x %d threads, %d B smem, %.0f ms\n",
                    e.k.name.c_str(), regs, e.k.ctas_per_sm, e.k.threads, e.k.smem_bytes, e.k.compile_ms);
    }

    void worker()
    {
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            cv_jobs.wait(lk, [&] { return stop || !jobs.empty(); });
            if (stop) return;
            std::shared_ptr<Entry> e = jobs.front();
            jobs.pop_front();
            lk.unlock();
            compile(*e);
            lk.lock();
            finish_compile(*e);
            ctr.pending--;
            cv_done.notify_all();
        }
    }

    void ensure_pool()
    {
        if (!pool.empty()) return;
        unsigned n = std::thread::hardware_concurrency();
        if (const char *s = std::getenv("B200SV_JIT_THREADS")) n = (unsigned)std::atoi(s);
        n = n < 1 ? 1 : (n > 16 ? 16 : n);
        for (unsigned i = 0; i < n; i++) pool.emplace_back([this] { worker(); });
    }
};

Engine::Engine() : p_(new Impl) {}

Engine::~Engine()
{
    {
        std::lock_guard<std::mutex> lk(p_->mu);
        p_->stop = true;
    }
    p_->cv_jobs.notify_all();
    for (auto &t : p_->pool) t.join();
    delete p_;
}

Engine &Engine::get()
{
    static Engine *e = new Engine; // never destroyed: the library is loaded RTLD_NODELETE by Catalyst
    return *e;
}

bool Engine::available(std::string *why)
{
    std::call_once(p_->once, [&] { load_api(p_->api); });
    if (why) *why = p_->api.why;
    return p_->api.ok;
}

const Kernel *Engine::lookup(int cuda_device, const std::string &key, int threshold, bool sync,
                             const std::function<std::string(Kernel &)> &gen)
{
    if (!available()) return nullptr;
    std::string full((const char *)&cuda_device, sizeof(int));
    full += key;
    std::unique_lock<std::mutex> lk(p_->mu);
    auto &slot = p_->cache[full];
    if (!slot) slot = std::make_shared<Entry>();
    std::shared_ptr<Entry> e = slot;
    if (e->state == READY) return &e->k;
    if (e->state == FAILED) return nullptr;
    if (e->state == COMPILED) {
        p_->make_ready(*e, cuda_device);
        return e->state == READY ? &e->k : nullptr;
    }
    if (e->state == COMPILING) {
        if (!sync) return nullptr;
        p_->cv_done.wait(lk, [&] { return e->state != COMPILING; });
        if (e->state == COMPILED) p_->make_ready(*e, cuda_device);
        return e->state == READY ? &e->k : nullptr;
    }
    if (++e->sightings < threshold) return nullptr;
    e->source = gen(e->k);
    e->state = COMPILING;
    if (sync) {
        lk.unlock();
        p_->compile(*e);
        lk.lock();
        p_->finish_compile(*e);
        if (e->state == COMPILED) p_->make_ready(*e, cuda_device);
        return e->state == READY ? &e->k : nullptr;
    }
    p_->ensure_pool();
    p_->ctr.pending++;
    p_->jobs.push_back(e);
    p_->cv_jobs.notify_one();
    return nullptr;
}

int Engine::launch(const Kernel &k, unsigned grid, double *psi, const double *m, void *stream)
{
    void *args[2] = {&psi, const_cast<double *>(m)};
    return p_->api.cuLaunchKernel(k.fn, grid, 1, 1, (unsigned)k.threads, 1, 1, (unsigned)k.smem_bytes, stream, args,
                                  nullptr);
}

void Engine::wait_all()
{
    std::unique_lock<std::mutex> lk(p_->mu);
    p_->cv_done.wait(lk, [&] { return p_->ctr.pending == 0; });
}

Counters Engine::counters()
{
    std::lock_guard<std::mutex> lk(p_->mu);
    return p_->ctr;
}

int Engine::sm_count(int cuda_device)
{
    std::lock_guard<std::mutex> lk(p_->mu);
    auto it = p_->sms.find(cuda_device);
    if (it != p_->sms.end()) return it->second;
    int n = 0;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, cuda_device);
    if (n <= 0) n = 148;
    p_->sms[cuda_device] = n;
    return n;
}

std::string Engine::source_of(const std::string &key)
{
    std::lock_guard<std::mutex> lk(p_->mu);
    for (auto &kv : p_->cache)
        if (kv.first.size() >= sizeof(int) && kv.first.compare(sizeof(int), std::string::npos, key) == 0)
            return kv.second->source;
    return std::string();
}

} // namespace jit
} // namespace b200
