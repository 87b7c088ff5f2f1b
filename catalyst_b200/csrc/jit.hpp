// jit.hpp — run-time specialisation of the fused tile pass.
//
// Why (profiles/README.md, r1_v3): the interpreted tile kernel (k_tile3) moves exactly the algorithmic
// bytes but is ISSUE-bound: ~270-320 thread instructions per amplitude per pass of which ~60 are FP64;
// 28 % of the issued instructions are per-op decode (control word, predicates, dispatch), 17 % the
// round prologue/epilogue (swizzle address arithmetic), 10 % constant-bank matrix loads.  All of that
// is a function of the pass STRUCTURE only (which bits, which masks, which matrix entries are 0/±1),
// not of the gate parameters.  So a pass whose structure recurs (every variational / shot loop, every
// bench step) is compiled once into a straight-line sm_100a kernel: bit positions, swizzle offsets
// and slot masks become immediates, matrix entries are read as direct constant-bank operands from the
// kernel parameter block (the values stay run-time arguments, so new angles reuse the kernel),
// permutation matrices (X, CNOT, SWAP) become register renames.
//
// NVRTC and the driver API are resolved with dlopen at first use (no link-time dependency: the
// library still loads on a box without a GPU driver).  Kernels are cached per process by their
// structural key; compilation runs on a small thread pool, and until a kernel is ready the
// interpreted k_tile3 path executes the pass (same results, lower speed).
//
// Compiled cubins also persist ACROSS processes: $B200SV_JIT_CACHE (default $XDG_CACHE_HOME/b200sv or
// ~/.cache/b200sv; "off" disables) holds one file per structural key (two 64-bit FNV-1a hashes of
// generator version + NVRTC version + key name the file; the full key is stored inside and compared),
// written atomically (temp file + rename).  A second process that meets the same pass structure loads
// the cubin at the FIRST sighting and never runs the interpreter or NVRTC for it.
// $B200SV_JIT_DUMP=<dir> additionally writes every generated source (<kernel>.cu) and cubin there.
#pragma once

#include <cstdint>
#include <functional>
#include <string>
#include <vector>

namespace b200 {
namespace jit {

struct Kernel {
    void *fn = nullptr;        // CUfunction
    int threads = 0;           // CTA size
    int smem_bytes = 0;        // dynamic shared memory per CTA
    int ctas_per_sm = 0;       // resident CTAs per SM (occupancy query)
    int num_regs = 0;
    int param_doubles = 0;     // doubles in the parameter block after the state pointer
    double fp64_per_thread = 0; // FP64 instructions one thread issues per tile (generator's count, upper bound)
    bool from_disk = false;    // loaded from the on-disk cubin cache instead of compiled
    double compile_ms = 0.0;
    std::string name;
    std::string error;         // non-empty: compilation failed, the pass stays on the interpreter
};

enum class Mode { Off = 0, Auto = 1, Sync = 2 };

struct Counters {
    int64_t compiled = 0, failed = 0, pending = 0, disk_hits = 0;
    double compile_ms = 0.0;
};

// Process-wide engine (kernels are shared by all device objects of the process, per CUDA device).
class Engine {
  public:
    static Engine &get();
    // false + reason when NVRTC or the driver API cannot be loaded
    bool available(std::string *why = nullptr);
    bool nvrtc_available(); // NVRTC alone (sources can be compiled to cubins without a driver)
    // development path: generate, compile to a cubin ($B200SV_JIT_DUMP, disk cache), never load or launch
    bool compile_offline(const std::string &key, const std::function<std::string(Kernel &)> &gen, Kernel &out,
                         std::string &source);
    // Look the structural key up.  Returns the kernel when it is ready.  Otherwise counts the
    // sighting and, once the key has been seen `threshold` times, compiles it: synchronously when
    // `sync`, else on the pool (nullptr is returned until it is ready).  `gen` produces the CUDA
    // source and fills threads / smem_bytes / param_doubles / name of the Kernel record.
    const Kernel *lookup(int cuda_device, const std::string &key, int threshold, bool sync,
                         const std::function<std::string(Kernel &)> &gen);
    // launch with parameter block = { psi pointer, m[param_doubles] }
    int launch(const Kernel &k, unsigned grid, double *psi, const double *m, void *stream);
    void wait_all(); // block until no compilation is pending
    // cancel queued compilations and wait for the running ones (last device of the process going away, exit)
    void quiesce();
    Counters counters();
    int sm_count(int cuda_device);
    // last generated source of a key (debugging / tests)
    std::string source_of(const std::string &key);

  private:
    Engine();
    ~Engine();
    struct Impl;
    Impl *p_;
};

} // namespace jit
} // namespace b200
