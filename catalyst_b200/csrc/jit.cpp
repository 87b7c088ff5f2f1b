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

#include <execinfo.h>
#include <fcntl.h>
#include <signal.h>
#include <sys/stat.h>
#include <sys/types.h>
#include <unistd.h>

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
    int (*nvrtcVersion)(int *, int *) = nullptr;
    std::string why;
    bool ok = false;       // NVRTC + driver API: kernels can be compiled, loaded and launched
    bool nvrtc_ok = false; // NVRTC alone: sources can be compiled to cubins (no driver needed)
    int nvrtc_major = 0, nvrtc_minor = 0;
};

template <typename F> bool sym(void *h, const char *name, F &out, std::string &why)
{
    out = reinterpret_cast<F>(dlsym(h, name));
    if (!out) why += std::string(" missing symbol ") + name + ";";
    return out != nullptr;
}

// $B200SV_DEBUG_SEGV=1 (diagnosis): print a native backtrace on SIGSEGV / SIGABRT, then leave with the usual code
void debug_fault_handler(int sig)
{
    void *frames[64];
    const int n = backtrace(frames, 64);
    const char msg[] = "[b200sv] fatal signal, native backtrace:\n";
    (void)!write(2, msg, sizeof msg - 1);
    backtrace_symbols_fd(frames, n, 2);
    _exit(128 + sig);
}

void load_api(Api &a)
{
    if (std::getenv("B200SV_DEBUG_SEGV")) {
        signal(SIGSEGV, debug_fault_handler);
        signal(SIGABRT, debug_fault_handler);
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
    ok &= sym(a.h_nvrtc, "nvrtcCreateProgram", a.nvrtcCreateProgram, a.why);
    ok &= sym(a.h_nvrtc, "nvrtcCompileProgram", a.nvrtcCompileProgram, a.why);
    ok &= sym(a.h_nvrtc, "nvrtcGetCUBINSize", a.nvrtcGetCUBINSize, a.why);
    ok &= sym(a.h_nvrtc, "nvrtcGetCUBIN", a.nvrtcGetCUBIN, a.why);
    ok &= sym(a.h_nvrtc, "nvrtcGetProgramLogSize", a.nvrtcGetProgramLogSize, a.why);
    ok &= sym(a.h_nvrtc, "nvrtcGetProgramLog", a.nvrtcGetProgramLog, a.why);
    ok &= sym(a.h_nvrtc, "nvrtcDestroyProgram", a.nvrtcDestroyProgram, a.why);
    ok &= sym(a.h_nvrtc, "nvrtcVersion", a.nvrtcVersion, a.why);
    if (!ok) return;
    a.nvrtcVersion(&a.nvrtc_major, &a.nvrtc_minor);
    a.nvrtc_ok = true;
    const char *cuda_names[] = {"libcuda.so.1", "libcuda.so"};
    for (const char *n : cuda_names)
        if ((a.h_cuda = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!a.h_cuda) {
        a.why = "libcuda.so.1 (driver API) not loadable";
        return;
    }
    ok &= sym(a.h_cuda, "cuInit", a.cuInit, a.why);
    ok &= sym(a.h_cuda, "cuModuleLoadData", a.cuModuleLoadData, a.why);
    ok &= sym(a.h_cuda, "cuModuleGetFunction", a.cuModuleGetFunction, a.why);
    ok &= sym(a.h_cuda, "cuFuncSetAttribute", a.cuFuncSetAttribute, a.why);
    ok &= sym(a.h_cuda, "cuFuncGetAttribute", a.cuFuncGetAttribute, a.why);
    ok &= sym(a.h_cuda, "cuOccupancyMaxActiveBlocksPerMultiprocessor", a.cuOccupancyMaxActiveBlocksPerMultiprocessor, a.why);
    ok &= sym(a.h_cuda, "cuLaunchKernel", a.cuLaunchKernel, a.why);
    ok &= sym(a.h_cuda, "cuGetErrorString", a.cuGetErrorString, a.why);
    if (ok && a.cuInit(0) != 0) {
        a.why = "cuInit failed";
        ok = false;
    }
    a.ok = ok;
}

// ---- on-disk cubin cache ---------------------------------------------------------------------------
uint64_t fnv1a(const std::string &s, uint64_t h)
{
    for (unsigned char c : s) {
        h ^= c;
        h *= 0x100000001b3ull;
    }
    return h;
}

std::string cache_dir()
{
    if (const char *e = std::getenv("B200SV_JIT_CACHE")) {
        const std::string v(e);
        return (v == "off" || v == "0" || v.empty()) ? std::string() : v;
    }
    if (const char *x = std::getenv("XDG_CACHE_HOME"))
        if (*x) return std::string(x) + "/b200sv";
    if (const char *h = std::getenv("HOME"))
        if (*h) return std::string(h) + "/.cache/b200sv";
    return std::string();
}

void mkdirs(const std::string &dir)
{
    for (size_t i = 1; i <= dir.size(); i++)
        if (i == dir.size() || dir[i] == '/') ::mkdir(dir.substr(0, i).c_str(), 0755);
}

template <typename T> void wr(std::string &o, const T &v) { o.append(reinterpret_cast<const char *>(&v), sizeof(T)); }
template <typename T> bool rd(const std::string &in, size_t &off, T &v)
{
    if (off + sizeof(T) > in.size()) return false;
    std::memcpy(&v, in.data() + off, sizeof(T));
    off += sizeof(T);
    return true;
}
bool rd_bytes(const std::string &in, size_t &off, size_t n, std::string &out)
{
    if (off + n > in.size()) return false;
    out.assign(in.data() + off, n);
    off += n;
    return true;
}

const uint32_t kCacheMagic = 0x42325356u; // "B2SV"
const uint32_t kCacheVersion = 2;
enum State { SEEN = 0, COMPILING = 1, COMPILED = 2, READY = 3, FAILED = 4 };

struct Entry {
    int state = SEEN;
    int sightings = 0;
    bool disk_checked = false;
    Kernel k;
    std::string key; // structural key (without the device ordinal): names the cache file
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

    // Ranks of one job (one process per GPU) meet the same pass structures at the same moment: the first
    // process to create <file>.lock compiles, the others wait for the cache file instead of running NVRTC
    // on the same source (8 ranks x 69 kernels was 84 s of CPU per rank).  A stale lock only costs the wait.
    void compile_shared(Entry &e)
    {
        const std::string path = cache_path(e.key);
        if (path.empty()) return compile(e);
        const std::string lock = path + ".lock";
        mkdirs(cache_dir());
        const int fd = ::open(lock.c_str(), O_CREAT | O_EXCL | O_WRONLY, 0644);
        if (fd >= 0) {
            ::close(fd);
            compile(e);
            unlink(lock.c_str());
            return;
        }
        const auto t0 = std::chrono::steady_clock::now();
        for (;;) {
            if (disk_load(e)) {
                e.k.error.clear();
                e.k.compile_ms = 0.0;
                return;
            }
            struct stat st;
            const bool locked = ::stat(lock.c_str(), &st) == 0;
            const double waited = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            if (!locked || waited > 90.0) break; // the compiling process gave up (or died): do it here
            std::this_thread::sleep_for(std::chrono::milliseconds(25));
        }
        if (disk_load(e)) {
            e.k.error.clear();
            return;
        }
        compile(e);
    }

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
        if (err.empty()) {
            disk_store(e);
            dump(e);
        }
    }

    std::string cache_path(const std::string &key) const
    {
        const std::string dir = cache_dir();
        if (dir.empty()) return std::string();
        const std::string salted = "nvrtc" + std::to_string(api.nvrtc_major) + "." + std::to_string(api.nvrtc_minor) + "/sm_100a/" + key;
        char name[64];
        snprintf(name, sizeof name, "/%016llx%016llx.b2sv", (unsigned long long)fnv1a(salted, 0xcbf29ce484222325ull),
                 (unsigned long long)fnv1a(salted, 0x84222325cbf29ce4ull));
        return dir + name;
    }

    void disk_store(const Entry &e) const // no lock needed: touches only e and the file system
    {
        const std::string path = cache_path(e.key);
        if (path.empty()) return;
        mkdirs(cache_dir());
        std::string blob;
        wr(blob, kCacheMagic);
        wr(blob, kCacheVersion);
        wr(blob, (uint64_t)e.key.size());
        blob += e.key;
        wr(blob, (uint64_t)e.k.name.size());
        blob += e.k.name;
        wr(blob, (int32_t)e.k.threads);
        wr(blob, (int32_t)e.k.smem_bytes);
        wr(blob, (int32_t)e.k.param_doubles);
        wr(blob, (double)e.k.fp64_per_thread);
        wr(blob, (uint64_t)e.cubin.size());
        blob.append(e.cubin.data(), e.cubin.size());
        char tmp[64];
        snprintf(tmp, sizeof tmp, ".tmp.%d.%p", (int)getpid(), (const void *)&e);
        const std::string tpath = path + tmp;
        FILE *f = fopen(tpath.c_str(), "wb");
        if (!f) return;
        const bool ok = fwrite(blob.data(), 1, blob.size(), f) == blob.size();
        fclose(f);
        if (!ok || rename(tpath.c_str(), path.c_str()) != 0) unlink(tpath.c_str());
    }

    bool disk_load(Entry &e) const
    {
        const std::string path = cache_path(e.key);
        if (path.empty()) return false;
        FILE *f = fopen(path.c_str(), "rb");
        if (!f) return false;
        std::string blob;
        char buf[1 << 16];
        size_t n;
        while ((n = fread(buf, 1, sizeof buf, f)) > 0) blob.append(buf, n);
        fclose(f);
        size_t off = 0;
        uint32_t magic = 0, ver = 0;
        uint64_t klen = 0, nlen = 0, clen = 0;
        int32_t threads = 0, smem = 0, pd = 0;
        double fp64 = 0;
        std::string key, name, cub;
        if (!rd(blob, off, magic) || !rd(blob, off, ver) || magic != kCacheMagic || ver != kCacheVersion) return false;
        if (!rd(blob, off, klen) || !rd_bytes(blob, off, (size_t)klen, key) || key != e.key) return false;
        if (!rd(blob, off, nlen) || !rd_bytes(blob, off, (size_t)nlen, name)) return false;
        if (!rd(blob, off, threads) || !rd(blob, off, smem) || !rd(blob, off, pd) || !rd(blob, off, fp64)) return false;
        if (!rd(blob, off, clen) || !rd_bytes(blob, off, (size_t)clen, cub) || cub.empty()) return false;
        e.k.name = name;
        e.k.threads = threads;
        e.k.smem_bytes = smem;
        e.k.param_doubles = pd;
        e.k.fp64_per_thread = fp64;
        e.k.from_disk = true;
        e.cubin.assign(cub.begin(), cub.end());
        return true;
    }

    void dump(const Entry &e) const
    {
        const char *dir = std::getenv("B200SV_JIT_DUMP");
        if (!dir || !*dir) return;
        mkdirs(dir);
        const std::string base = std::string(dir) + "/" + e.k.name;
        if (FILE *f = fopen((base + ".cu").c_str(), "w")) {
            fwrite(e.source.data(), 1, e.source.size(), f);
            fclose(f);
        }
        if (FILE *f = fopen((base + ".cubin").c_str(), "wb")) {
            fwrite(e.cubin.data(), 1, e.cubin.size(), f);
            fclose(f);
        }
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
        if (e.k.from_disk) ctr.disk_hits++;
        else ctr.compiled++;
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
            active++;
            lk.unlock();
            compile_shared(*e);
            lk.lock();
            finish_compile(*e);
            active--;
            ctr.pending--;
            {
                static std::once_flag late;
                std::call_once(late, [] { std::atexit(&Impl::quiesce_pool_at_exit); });
            }
            cv_done.notify_all();
        }
    }

    // The engine is never destroyed (the library may be loaded RTLD_NODELETE), but a process must not run static
    // destructors (NVRTC's among them) while a pool thread is still inside nvrtcCompileProgram: at exit queued
    // jobs are dropped and the handler waits (bounded) until no compilation is in flight.  The threads themselves
    // are left parked on their condition variable; the process is going away.
    int active = 0; // compilations running on pool threads
    // Drop queued compilations (their entries go back to "seen", so a later device resubmits them) and wait,
    // bounded, until no pool thread is inside NVRTC.  Called when the LAST device object of the process is
    // destroyed and again from atexit: a process must not run exit-time destructors (NVRTC registers some
    // lazily, during its first compilation, i.e. AFTER any handler registered here) while a pool thread is
    // still inside nvrtcCompileProgram -- seen on the GPU box as SIGSEGV at exit of a short-lived process
    // whose last call (State() restoring the wire order) had just submitted a new pass structure.
    void quiesce(int max_seconds)
    {
        std::unique_lock<std::mutex> lk(mu);
        for (auto &e : jobs) {
            e->state = SEEN;
            ctr.pending--;
        }
        jobs.clear();
        cv_done.notify_all();
        cv_done.wait_for(lk, std::chrono::seconds(max_seconds), [&] { return active == 0; });
    }
    static void quiesce_pool_at_exit() { Engine::get().p_->quiesce(30); }

    void ensure_pool()
    {
        if (!pool.empty()) return;
        static std::once_flag reg;
        std::call_once(reg, [] { std::atexit(&Impl::quiesce_pool_at_exit); });
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
    if (!e->disk_checked) { // first sighting in this process: another process may have compiled this structure
        e->disk_checked = true;
        e->key = key;
        if (p_->disk_load(*e)) {
            e->state = COMPILED;
            p_->make_ready(*e, cuda_device);
            return e->state == READY ? &e->k : nullptr;
        }
    }
    if (++e->sightings < threshold) return nullptr;
    e->source = gen(e->k);
    e->state = COMPILING;
    if (sync) {
        lk.unlock();
        p_->compile(*e);
        lk.lock();
        p_->finish_compile(*e);
        // rank threads of one process share kernels (the key does not hold the rank): a peer that met this key
        // while it was being compiled here waits on cv_done
        p_->cv_done.notify_all();
        if (e->state == COMPILED) p_->make_ready(*e, cuda_device);
        return e->state == READY ? &e->k : nullptr;
    }
    p_->ensure_pool();
    p_->ctr.pending++;
    p_->jobs.push_back(e);
    p_->cv_jobs.notify_one();
    return nullptr;
}

bool Engine::nvrtc_available()
{
    std::call_once(p_->once, [&] { load_api(p_->api); });
    return p_->api.nvrtc_ok;
}

// Offline path (development, no GPU needed): generate + compile + dump, never loaded or launched.
bool Engine::compile_offline(const std::string &key, const std::function<std::string(Kernel &)> &gen, Kernel &out,
                             std::string &source)
{
    if (!nvrtc_available()) return false;
    Entry e;
    e.key = key;
    e.source = gen(e.k);
    p_->compile(e);
    out = e.k;
    source = e.source;
    return e.k.error.empty();
}

int Engine::launch(const Kernel &k, unsigned grid, double *psi, const double *m, void *stream)
{
    void *args[2] = {&psi, const_cast<double *>(m)};
    return p_->api.cuLaunchKernel(k.fn, grid, 1, 1, (unsigned)k.threads, 1, 1, (unsigned)k.smem_bytes, stream, args,
                                  nullptr);
}

void Engine::quiesce() { p_->quiesce(60); }

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
