// simulator_jit.cpp — plans one tile pass for the specialised (run-time compiled) kernel and emits its
// CUDA source.  Same tile geometry, commutation rule and round structure as launch_pass3
// (simulator.cpp); what differs is that a thread owns 2^NR amplitudes (NR = 4 by default instead of
// 3, the interpreter's limit), every op whose non-diagonal bits fit the register bits runs in
// registers (incl. 4-bit PAIRs such as DoubleExcitation and general Pauli rotations), and all
// structure is folded into the instruction stream (jit.hpp explains why).
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <sstream>

#include "jit.hpp"
#include "simulator.hpp"

namespace b200 {

#define B200_FAIL_IF(cond, msg)                                                                    \
    do {                                                                                           \
        if (cond) throw ::b200::Error(std::string(msg));                                           \
    } while (0)

namespace {

enum { CLS_ZERO = 0, CLS_ONE = 1, CLS_MONE = 2, CLS_GEN = 3 };

struct JOp {
    int kind = 0;               // GateOp::Kind
    uint32_t x = 0, a = 0, b = 0; // tile-local target mask / patterns
    uint32_t cin_m = 0, cin_v = 0; // control predicate inside the tile
    uint64_t co_m = 0, co_v = 0;   // control predicate outside the tile (incl. rank bits)
    uint32_t zin = 0;              // z mask inside the tile
    uint64_t zo = 0;               // z mask outside the tile
    int ny = 0;
    int nm = 0, moff = 0;          // doubles in the parameter block
    uint32_t cls = 0;              // 2 bits per double
    bool literal = false;          // entries may be classified (0 / +-1 / general) at generation time
    double vals[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    std::vector<int> tg;           // DENSE: tile-local target positions, most significant first
    std::vector<double> dvals;     // DENSE: 2 * 4^k doubles, row-major (re, im)
    int uacc = -1;                 // >= 0: tile-uniform diagonal op, folded into uniform accumulator `uacc`
};

// A tile-uniform accumulator: the product of the diagonal ops of one fused run (one register-bit signature)
// whose predicate / parity depends on bits OUTSIDE the tile only.  Its value is the same for every thread of
// the CTA, so ONE warp evaluates it per tile (one op per lane, warp-shuffle product tree) instead of every
// thread multiplying through the whole fan (a QFT fan: 19 of its 29 controlled phases).
struct UAcc {
    int type = 0;   // 0 PHASE (one complex), 1 PARITY (even / odd pair)
    int n = 0;      // ops
    int first = 0;  // first table entry (entries of one accumulator are contiguous)
};

struct JRound {
    uint32_t S = 0; // tile-local register bits
    std::vector<JOp> ops;
};

struct JPlan {
    int t = 0, c = 0, nr = 0, nl = 0, dbuf = 0, fuse_diag = 1;
    std::vector<int> high;
    uint64_t rank_base = 0;
    std::vector<JRound> rounds;
    int nm = 0;
    std::vector<UAcc> uaccs;
    int nu = 0, uoff = 0; // uniform-table entries (8 doubles each) and their offset in the parameter doubles
    int direct_store = 0; // option: the last round stores its registers straight to global memory
    int memhint = 0;      // option: cache hints (Options::jit_memhint)
    int uniform = 1;      // option: evaluate tile-uniform diagonal factors once per tile
};

inline int cls_of(double v) { return v == 0.0 ? CLS_ZERO : (v == 1.0 ? CLS_ONE : (v == -1.0 ? CLS_MONE : CLS_GEN)); }

// Scale factoring.  An uncontrolled one-qubit gate U = u00 * U' with U' = U / u00 has a literal 1 on its
// diagonal, which removes a quarter (real rotations: half) of its FP64 work: RX/RY cost one FMA per output
// component instead of a multiply and an FMA, Hadamard becomes an add and a subtract.  The scalars u00
// commute with every op of the pass, so their product is folded into one op that multiplies every amplitude
// anyway: an uncontrolled parity phase (RZ, IsingZZ, MultiRZ) if the pass holds one, else the last factored
// gate (which then keeps its general form).  Which ops are factored depends only on the structure of the
// pass (exact zeros on the diagonal) except for |u00| < 1e-4, where the op is left alone: the passes of the
// 30-qubit circuits are bound by the FP64 pipe, so this is worth 15-25% of their time.
inline void factor_scales(JPlan &P)
{
    using cd = std::complex<double>;
    cd total(1.0, 0.0);
    JOp *last = nullptr, *absorber = nullptr;
    for (JRound &R : P.rounds)
        for (JOp &o : R.ops) {
            const bool everywhere = o.cin_m == 0 && o.co_m == 0;
            if (o.kind == (int)GateOp::PARITY && everywhere && !absorber) absorber = &o;
            if (o.kind != (int)GateOp::PAIR || !everywhere || __builtin_popcount(o.x) != 1 || o.a != 0 || o.b != o.x)
                continue;
            const cd u00(o.vals[0], o.vals[1]);
            if (std::abs(u00) < 1e-4) continue;
            if (u00 == cd(1.0, 0.0)) continue;
            for (int q = 1; q < 4; q++) {
                const cd v = cd(o.vals[2 * q], o.vals[2 * q + 1]) / u00;
                o.vals[2 * q] = v.real();
                o.vals[2 * q + 1] = v.imag();
            }
            o.vals[0] = 1.0;
            o.vals[1] = 0.0;
            total *= u00;
            last = &o;
        }
    if (!last) return;
    if (absorber) {
        for (int q = 0; q < 2; q++) {
            const cd v = cd(absorber->vals[2 * q], absorber->vals[2 * q + 1]) * total;
            absorber->vals[2 * q] = v.real();
            absorber->vals[2 * q + 1] = v.imag();
        }
        return;
    }
    for (int q = 0; q < 4; q++) {
        const cd v = cd(last->vals[2 * q], last->vals[2 * q + 1]) * total;
        last->vals[2 * q] = v.real();
        last->vals[2 * q + 1] = v.imag();
    }
}

inline uint32_t fold3(uint32_t i) { return ((i >> 3) ^ (i >> 6) ^ (i >> 9) ^ (i >> 12)) & 7u; }
inline uint32_t swz(uint32_t i) { return i ^ fold3(i); }

template <typename T> void put(std::string &s, const T &v) { s.append(reinterpret_cast<const char *>(&v), sizeof(T)); }

std::string hex32(uint32_t v)
{
    char b[16];
    snprintf(b, sizeof b, "0x%xu", v);
    return b;
}
std::string hex64(uint64_t v)
{
    char b[32];
    snprintf(b, sizeof b, "0x%llxull", (unsigned long long)v);
    return b;
}

// ---- expression builder: sum of sign * coef * operand with coefficient classes ---------------------
struct Term {
    int cls;
    std::string coef; // expression of a GEN coefficient
    std::string opnd;
    bool neg;
};

// FP64-pipe instructions emitted so far by the current generate() call (DFMA / DMUL / DADD each count one)
thread_local double g_fp64_emitted = 0.0;

std::string lin(const std::vector<Term> &terms)
{
    std::string acc;
    // operands with coefficient +-1 first: they start the FMA chain for free
    std::vector<const Term *> ord;
    for (auto &t : terms)
        if (t.cls == CLS_ONE || t.cls == CLS_MONE) ord.push_back(&t);
    for (auto &t : terms)
        if (t.cls == CLS_GEN) ord.push_back(&t);
    for (const Term *t : ord) {
        bool neg = t->neg;
        if (t->cls == CLS_MONE) neg = !neg;
        g_fp64_emitted += (t->cls != CLS_GEN && acc.empty()) ? 0.0 : 1.0;
        if (t->cls != CLS_GEN) {
            if (acc.empty()) acc = neg ? "(-" + t->opnd + ")" : t->opnd;
            else acc = "(" + acc + (neg ? " - " : " + ") + t->opnd + ")";
        }
        else {
            if (acc.empty()) acc = neg ? "(-" + t->coef + " * " + t->opnd + ")" : "(" + t->coef + " * " + t->opnd + ")";
            else acc = "fma(" + std::string(neg ? "-" : "") + t->coef + ", " + t->opnd + ", " + acc + ")";
        }
    }
    return acc.empty() ? "0.0" : acc;
}

struct Coef { // a complex coefficient: literal classes + expressions
    int cr, ci;
    std::string er, ei;
};

// (re, im) of coef * (xr, xi) [+ coef2 * (yr, yi)]
void cmul_terms(std::vector<Term> &re, std::vector<Term> &im, const Coef &k, const std::string &xr, const std::string &xi)
{
    re.push_back({k.cr, k.er, xr, false});
    re.push_back({k.ci, k.ei, xi, true});
    im.push_back({k.cr, k.er, xi, false});
    im.push_back({k.ci, k.ei, xr, false});
}

// Marks the diagonal ops that fold into tile-uniform accumulators (see UAcc).  Mirrors the run / accumulator
// formation of emit_diag_run in generate(): a run is a maximal sequence of consecutive PHASE / PARITY ops of
// a round (length >= 2); inside a run an accumulator is identified by its register-bit signature.
inline bool diag_kind(int k) { return k == (int)GateOp::PHASE || k == (int)GateOp::PARITY; }

void analyze_uniform(JPlan &P)
{
    P.uaccs.clear();
    P.nu = 0;
    const int threads = 1 << (P.t - P.nr);
    if (!P.uniform || !P.fuse_diag || threads < 32) return;
    for (JRound &R : P.rounds) {
        for (size_t i0 = 0; i0 < R.ops.size();) {
            if (!diag_kind(R.ops[i0].kind)) {
                i0++;
                continue;
            }
            size_t i1 = i0;
            while (i1 < R.ops.size() && diag_kind(R.ops[i1].kind)) i1++;
            if (i1 - i0 >= 2) {
                struct Sig {
                    int type;
                    uint32_t rcm, rcv, rz;
                    std::vector<size_t> idx;
                };
                std::vector<Sig> sigs;
                for (size_t i = i0; i < i1; i++) {
                    const JOp &g = R.ops[i];
                    const bool parity = g.kind == (int)GateOp::PARITY;
                    if (g.cin_m & ~R.S) continue;                       // thread-bit predicate: per thread
                    if (parity && (g.zin & ~R.S)) continue;             // thread-bit parity: per thread
                    if (!g.co_m && !(parity && g.zo)) continue;         // nothing outside: a plain constant
                    const uint32_t rcm = g.cin_m & R.S, rcv = g.cin_v & R.S, rz = parity ? (g.zin & R.S) : 0u;
                    Sig *hit = nullptr;
                    for (auto &sg : sigs)
                        if (sg.type == (parity ? 1 : 0) && sg.rcm == rcm && sg.rcv == rcv && sg.rz == rz) hit = &sg;
                    if (!hit) {
                        sigs.push_back(Sig{parity ? 1 : 0, rcm, rcv, rz, {}});
                        hit = &sigs.back();
                    }
                    hit->idx.push_back(i);
                }
                for (auto &sg : sigs) {
                    if (sg.idx.size() < 3) continue; // a warp-collective per tile must replace enough per-thread work
                    UAcc u;
                    u.type = sg.type;
                    u.n = (int)sg.idx.size();
                    u.first = P.nu;
                    for (size_t i : sg.idx) R.ops[i].uacc = (int)P.uaccs.size();
                    P.nu += u.n;
                    P.uaccs.push_back(u);
                }
            }
            i0 = i1;
        }
    }
}

// ---- CUDA source of one pass ---------------------------------------------------------------------
std::string generate(const JPlan &P, jit::Kernel &K, const std::string &name)
{
    const int t = P.t, c = P.c, nr = P.nr;
    const int tbits = t - nr;          // log2(threads)
    const uint32_t TT = 1u << tbits;
    const uint32_t nelem = 1u << t;
    const uint32_t niter = nelem / TT; // = 2^nr
    const uint32_t nslots = 1u << nr;
    const uint64_t ntiles = 1ull << (P.nl - t);
    std::ostringstream o;
    g_fp64_emitted = 0.0;
    auto mref = [&](int k) { return "P.m[" + std::to_string(k) + "]"; };
    // global offset of tile-local index e (an OR over disjoint bit sets)
    auto off_of = [&](uint32_t e) -> uint64_t {
        uint64_t r = e & ((1u << c) - 1u);
        for (size_t h = 0; h < P.high.size(); h++)
            if ((e >> (c + h)) & 1u) r |= 1ull << P.high[h];
        return r;
    };

    o << "// generated by catalyst_b200/csrc/simulator_jit.cpp: t=" << t << " c=" << c << " nr=" << nr
      << " rounds=" << P.rounds.size() << "\n";
    o << "typedef double2 cplx;\n";
    // rank_base (rank << local bits) is a run-time argument: every rank of a sharded job runs the same cubin
    o << "struct Params { unsigned long long rank_base; double m[" << std::max(1, P.nm) << "]; };\n";
    o << "__device__ __forceinline__ void cpa16(void *d, const void *s) {\n"
         "  const unsigned a = (unsigned)__cvta_generic_to_shared(d);\n"
         "  asm volatile(\"cp.async.cg.shared.global" << std::string((P.memhint & 2) && c >= 4 ? ".L2::256B" : (P.memhint & 4) ? ".L2::128B" : "")
      << " [%0], [%1], 16;\\n\" :: \"r\"(a), \"l\"(s) : \"memory\");\n}\n";
    o << "__device__ __forceinline__ unsigned long long base_of(unsigned long long tile) {\n"
         "  unsigned long long b = tile << "
      << c << ";\n";
    for (size_t h = 0; h < P.high.size(); h++) {
        const int p = P.high[h];
        // insert a zero at position p unless the positions so far are contiguous from c (then the
        // shift by c already made room)
        o << "  b = ((b >> " << p << ") << " << (p + 1) << ") | (b & " << hex64((1ull << p) - 1ull) << ");\n";
    }
    o << "  return b;\n}\n";
    const int usmem = P.uaccs.empty() ? 0 : 8 * (8 * P.nu + 4 * (int)P.uaccs.size());
    const int minb = P.dbuf ? std::max(1, std::min(8, (227 * 1024) / (int)(2 * 16 * nelem + 1024 + usmem)))
                            : std::max(1, std::min(8, (227 * 1024) / (int)(16 * nelem + 1024 + usmem)));
    // register budget: leave the compiler room for 2^nr amplitudes (4 registers each) + temporaries
    int minb_regs = (int)(65536 / (TT * (uint32_t)(4 * nslots + 56)));
    // (measured at 30 qubits, 16 amplitudes per thread: 3 CTAs x 128 threads with the cp.async double
    // buffer 137 ms per circuit; 4 CTAs without it 150 ms; 5 CTAs, 102 registers, spilling: 190 ms)
    const int launch_minb = std::max(1, std::min(minb, std::max(1, minb_regs)));
    if (!P.uaccs.empty()) {
        o << "__constant__ unsigned UMETA[" << P.uaccs.size() << "] = {";
        for (size_t u = 0; u < P.uaccs.size(); u++)
            o << (u ? ", " : "") << (((uint32_t)P.uaccs[u].n << 16) | (uint32_t)P.uaccs[u].first) << "u";
        o << "};\n";
    }
    o << "extern \"C\" __global__ void __launch_bounds__(" << TT << ", " << launch_minb << ") " << name
      << "(cplx *__restrict__ psi, const __grid_constant__ Params P)\n{\n";
    o << "  extern __shared__ __align__(16) cplx smem[];\n";
    o << "  const unsigned tid = threadIdx.x;\n";
    const uint32_t tile_bufs = P.dbuf ? 2u : 1u;
    if (!P.uaccs.empty()) {
        // tile-uniform diagonal factors: op table (8 doubles per op: predicate mask / value, z mask, e0, e1) copied
        // once from the parameter block, and one (even, odd) result pair per accumulator, rewritten every tile
        o << "  double *const utab = reinterpret_cast<double *>(smem + " << tile_bufs * nelem << "u);\n";
        o << "  double *const su = utab + " << 8 * P.nu << "u;\n";
        o << "  for (unsigned i = tid; i < " << 8 * P.nu << "u; i += " << TT << "u) utab[i] = P.m[" << P.uoff << "u + i];\n";
        o << "  __syncthreads();\n";
    }
    // off_tid: tid covers tile-local bits 0..tbits-1
    {
        o << "  const unsigned long long off_tid = (unsigned long long)(tid & " << hex32((1u << std::min(c, tbits)) - 1u) << ")";
        for (int bpos = c; bpos < tbits; bpos++)
            o << " | ((unsigned long long)((tid >> " << bpos << ") & 1u) << " << P.high[bpos - c] << ")";
        o << ";\n";
    }
    o << "  const unsigned swtid = tid ^ (((tid >> 3) ^ (tid >> 6) ^ (tid >> 9) ^ (tid >> 12)) & 7u);\n";
    auto slot_expr = [&](const std::string &sw, uint32_t pat) { // smem index of (thread base | pattern)
        const uint32_t k = swz(pat), kl = k & 7u, kh = k & ~7u;
        std::string e = kl ? "(" + sw + " ^ " + std::to_string(kl) + "u)" : sw;
        if (kh) e = "(" + e + " + " + std::to_string(kh) + "u)";
        return e;
    };
    if (P.dbuf) {
        o << "  auto prefetch = [&](unsigned long long tile, cplx *buf) {\n"
             "    const cplx *src = psi + (base_of(tile) | off_tid);\n";
        for (uint32_t k = 0; k < niter; k++)
            o << "    cpa16(&buf[" << slot_expr("swtid", k * TT) << "], src + " << hex64(off_of(k * TT)) << ");\n";
        o << "    asm volatile(\"cp.async.commit_group;\\n\" ::: \"memory\");\n  };\n";
        o << "  unsigned cur = 0;\n";
        o << "  if (blockIdx.x < " << hex64(ntiles) << ") prefetch(blockIdx.x, smem);\n";
    }
    o << "  for (unsigned long long tile = blockIdx.x; tile < " << hex64(ntiles) << "; tile += gridDim.x) {\n";
    o << "    const unsigned long long base = base_of(tile);\n";
    o << "    const unsigned long long obase = base | P.rank_base;\n";
    o << "    cplx *__restrict__ tp = psi + (base | off_tid);\n";
    if (P.dbuf) {
        o << "    cplx *__restrict__ s = smem + cur * " << nelem << "u;\n";
        o << "    const bool has_next = tile + gridDim.x < " << hex64(ntiles) << ";\n";
        o << "    if (has_next) prefetch(tile + gridDim.x, smem + (cur ^ 1u) * " << nelem << "u);\n";
    }
    else {
        o << "    cplx *__restrict__ s = smem;\n";
        for (uint32_t k = 0; k < niter; k++)
            o << "    s[" << slot_expr("swtid", k * TT) << "] = tp[" << hex64(off_of(k * TT)) << "];\n";
    }
    if (!P.uaccs.empty()) {
        // Tile-uniform factors.  Accumulator g is evaluated by a group of four lanes: each lane multiplies through
        // every fourth op of the accumulator (table in shared memory), two shuffle steps combine the group, its
        // first lane publishes the (even, odd) pair.  Every warp runs the same straight code (idle groups see
        // n = 0), so the shuffles sit in convergent code.
        const uint32_t nwarps = TT / 32u, per_iter = 8u * nwarps, nacc = (uint32_t)P.uaccs.size();
        bool any_par = false;
        for (const UAcc &A : P.uaccs) any_par = any_par || A.type == 1;
        o << "    { // tile-uniform diagonal factors: " << nacc << " accumulators, " << P.nu << " ops\n";
        o << "      const unsigned lane = tid & 31u, grp = (tid >> 5) * 8u + (lane >> 2);\n";
        for (uint32_t base = 0; base < nacc; base += per_iter) {
            o << "      {\n";
            o << "        const unsigned acc = " << base << "u + grp;\n";
            o << "        const unsigned meta = acc < " << nacc << "u ? UMETA[acc] : 0u;\n";
            o << "        const unsigned first = meta & 0xffffu, n = meta >> 16;\n";
            o << "        double ar = 1.0, ai = 0.0" << (any_par ? ", br = 1.0, bi = 0.0" : "") << ";\n";
            o << "        for (unsigned k = lane & 3u; k < n; k += 4u) {\n";
            o << "          const double2 *e = reinterpret_cast<const double2 *>(utab) + 4u * (first + k);\n";
            o << "          const double2 mk = e[0];\n";
            o << "          if ((obase & (unsigned long long)__double_as_longlong(mk.x)) == (unsigned long long)__double_as_longlong(mk.y)) {\n";
            if (any_par) {
                o << "            const double2 zz = e[1], e0 = e[2], e1 = e[3];\n";
                o << "            const bool p = (__popcll(obase & (unsigned long long)__double_as_longlong(zz.x)) & 1) != 0;\n";
                o << "            const double xr = p ? e1.x : e0.x, xi = p ? e1.y : e0.y, yr = p ? e0.x : e1.x, yi = p ? e0.y : e1.y;\n";
                o << "            { const double tr = ar; ar = fma(-ai, xi, tr * xr); ai = fma(ai, xr, tr * xi); }\n";
                o << "            { const double tr = br; br = fma(-bi, yi, tr * yr); bi = fma(bi, yr, tr * yi); }\n";
            }
            else {
                o << "            const double2 e0 = e[2];\n";
                o << "            { const double tr = ar; ar = fma(-ai, e0.y, tr * e0.x); ai = fma(ai, e0.x, tr * e0.y); }\n";
            }
            o << "          }\n        }\n";
            for (int d = 1; d <= 2; d <<= 1) {
                o << "        { const double xr = __shfl_xor_sync(0xffffffffu, ar, " << d << "), xi = __shfl_xor_sync(0xffffffffu, ai, " << d << ");\n"
                     "          const double tr = ar; ar = fma(-ai, xi, tr * xr); ai = fma(ai, xr, tr * xi); }\n";
                if (any_par)
                    o << "        { const double xr = __shfl_xor_sync(0xffffffffu, br, " << d << "), xi = __shfl_xor_sync(0xffffffffu, bi, " << d << ");\n"
                         "          const double tr = br; br = fma(-bi, xi, tr * xr); bi = fma(bi, xr, tr * xi); }\n";
            }
            o << "        if ((lane & 3u) == 0u && acc < " << nacc << "u) {\n";
            o << "          reinterpret_cast<double2 *>(su)[2u * acc] = make_double2(ar, ai);\n";
            if (any_par) o << "          reinterpret_cast<double2 *>(su)[2u * acc + 1u] = make_double2(br, bi);\n";
            o << "        }\n      }\n";
        }
        o << "    }\n";
    }
    if (P.dbuf)
        o << "    if (has_next) asm volatile(\"cp.async.wait_group 1;\\n\" ::: \"memory\");\n"
             "    else asm volatile(\"cp.async.wait_group 0;\\n\" ::: \"memory\");\n";
    o << "    __syncthreads();\n";


    // ---- rounds ------------------------------------------------------------------------------------
    int rix = 0;
    bool direct_store = false;
    for (const JRound &R : P.rounds) {
        std::vector<int> rb, q;
        for (int b = 0; b < t; b++) ((R.S >> b) & 1u ? rb : q).push_back(b);
        // thread bit k -> tile-local position qq[k]; the three lowest thread bits go to positions that
        // are distinct mod 3 where possible, so that the 8 threads of a quarter warp hit 8 different
        // 16-byte bank groups under the XOR-fold swizzle
        std::vector<int> qq;
        {
            std::vector<char> used(q.size(), 0);
            bool res[3] = {false, false, false};
            for (size_t i = 0; i < q.size() && qq.size() < 3; i++)
                if (!res[q[i] % 3]) {
                    res[q[i] % 3] = true;
                    used[i] = 1;
                    qq.push_back(q[i]);
                }
            std::sort(qq.begin(), qq.end());
            for (size_t i = 0; i < q.size(); i++)
                if (!used[i]) qq.push_back(q[i]);
        }
        std::vector<uint32_t> pat(nslots, 0);
        for (uint32_t k = 0; k < nslots; k++)
            for (int j = 0; j < nr; j++)
                if ((k >> j) & 1u) pat[k] |= 1u << rb[j];
        auto slot_of = [&](uint32_t p) -> uint32_t { // register pattern (subset of S) -> slot number
            uint32_t k = 0;
            for (int j = 0; j < nr; j++)
                if ((p >> rb[j]) & 1u) k |= 1u << j;
            return k;
        };
        o << "    { // round " << rix++ << ": register bits";
        for (int b : rb) o << " " << b;
        o << "\n";
        // ib: deposit the thread bits (runs of consecutive positions move together)
        o << "      const unsigned ib = 0u";
        for (int k = 0; k < (int)qq.size();) {
            int len = 1;
            while (k + len < (int)qq.size() && qq[k + len] == qq[k] + len) len++;
            const uint32_t m = ((1u << len) - 1u) << k;
            const int sh = qq[k] - k;
            o << " | ((tid & " << hex32(m) << ")" << (sh > 0 ? " << " + std::to_string(sh) : (sh < 0 ? " >> " + std::to_string(-sh) : "")) << ")";
            k += len;
        }
        o << ";\n";
        o << "      const unsigned sb = ib ^ (((ib >> 3) ^ (ib >> 6) ^ (ib >> 9) ^ (ib >> 12)) & 7u);\n";
        for (uint32_t k = 0; k < nslots; k++) {
            o << "      double v" << k << "r, v" << k << "i; { const cplx q_ = s[" << slot_expr("sb", pat[k]) << "]; v" << k
              << "r = q_.x; v" << k << "i = q_.y; }\n";
        }
        auto V = [&](uint32_t k, char part) { return "v" + std::to_string(k) + part; };
        // ---- a run of consecutive diagonal ops (PHASE / PARITY) is fused: each op multiplies a
        // per-thread accumulator (one per distinct register-bit signature) instead of its 4..2^nr
        // register slots, run-time conditions become selects (no branches), and every slot is
        // multiplied once by the product of the accumulators that apply to it.  The QFT's fans of
        // controlled phases and QAOA's ZZ layers collapse from ~8 complex multiplies per op to ~1.
        auto cond_of = [&](const JOp &g) {
            std::string cond;
            if (g.co_m) cond = "((obase & " + hex64(g.co_m) + ") == " + hex64(g.co_v) + ")";
            const uint32_t tcm = g.cin_m & ~R.S, tcv = g.cin_v & ~R.S;
            if (tcm) cond += std::string(cond.empty() ? "" : " && ") + "((ib & " + hex32(tcm) + ") == " + hex32(tcv) + ")";
            return cond;
        };
        auto emit_diag_run = [&](size_t i0, size_t i1) {
            struct Acc {
                int type; // 0 PHASE, 1 PARITY
                uint32_t rcm, rcv, rz;
                Coef a, b; // current value (literal coefficient or variable)
                int id;
            };
            std::vector<Acc> accs;
            int uid = 0;
            o << "      { // fused run of " << (i1 - i0) << " diagonal ops\n";
            auto mul_into = [&](Coef &acc, const Coef &f, const std::string &name) {
                // acc = acc * f, result in fresh variables name{r,i}
                std::vector<Term> re, im;
                cmul_terms(re, im, f, acc.er, acc.ei);
                // cmul_terms treats f as the coefficient with classes and acc as operands: acc must be a variable or
                // GEN literal expression; classes of acc are ignored (operands), which is correct but not minimal
                o << "        const double " << name << "r = " << lin(re) << ", " << name << "i = " << lin(im) << ";\n";
                acc = Coef{CLS_GEN, CLS_GEN, name + "r", name + "i"};
            };
            // acc = acc * f only where `cond` holds: predicated FP64 instructions with the constants as direct
            // operands (a select-then-multiply costs 4 FSEL + 2..4 LDC more per op)
            auto mul_into_if = [&](Coef &acc, const Coef &f, const std::string &cond, const std::string &name) {
                std::vector<Term> re, im;
                cmul_terms(re, im, f, acc.er, acc.ei);
                o << "        double " << name << "r = " << acc.er << ", " << name << "i = " << acc.ei << ";\n";
                o << "        if (" << cond << ") { " << name << "r = " << lin(re) << "; " << name << "i = " << lin(im) << "; }\n";
                acc = Coef{CLS_GEN, CLS_GEN, name + "r", name + "i"};
            };
            std::vector<char> useen(P.uaccs.size(), 0);
            for (size_t i = i0; i < i1; i++) {
                const JOp &g = R.ops[i];
                std::string cond = cond_of(g);
                const uint32_t rcm = g.cin_m & R.S, rcv = g.cin_v & R.S;
                const int u = uid++;
                const std::string un = "f" + std::to_string(u);
                const bool uni = g.uacc >= 0;
                if (uni) { // tile-uniform op: its whole accumulator was evaluated once per tile (su[])
                    if (useen[(size_t)g.uacc]) continue;
                    useen[(size_t)g.uacc] = 1;
                    cond.clear();
                }
                if (g.kind == GateOp::PHASE) {
                    Coef f{(int)(g.cls & 3u), (int)((g.cls >> 2) & 3u), mref(g.moff), mref(g.moff + 1)};
                    if (uni) {
                        o << "        const double " << un << "r = su[" << 4 * g.uacc << "], " << un << "i = su[" << 4 * g.uacc + 1 << "];\n";
                        f = Coef{CLS_GEN, CLS_GEN, un + "r", un + "i"};
                    }
                    Acc *A = nullptr;
                    for (auto &a : accs)
                        if (a.type == 0 && a.rcm == rcm && a.rcv == rcv) A = &a;
                    if (A && !cond.empty()) {
                        mul_into_if(A->a, f, cond, "g" + std::to_string(A->id) + "_" + std::to_string(u));
                        continue;
                    }
                    if (!cond.empty()) {
                        o << "        const bool " << un << "c = " << cond << ";\n";
                        o << "        const double " << un << "r = " << un << "c ? " << mref(g.moff) << " : 1.0, " << un << "i = " << un
                          << "c ? " << mref(g.moff + 1) << " : 0.0;\n";
                        f = Coef{CLS_GEN, CLS_GEN, un + "r", un + "i"};
                    }
                    if (!A) {
                        accs.push_back(Acc{0, rcm, rcv, 0, f, f, (int)accs.size()});
                        continue;
                    }
                    if (A->a.cr != CLS_GEN || A->a.ci != CLS_GEN) std::swap(A->a, f); // keep the literal as the coefficient
                    mul_into(A->a, f, "g" + std::to_string(A->id) + "_" + std::to_string(u));
                }
                else { // PARITY: factor e0 on even parity of the z bits, e1 on odd
                    const uint32_t tz = g.zin & ~R.S, rz = g.zin & R.S;
                    Coef fa{(int)(g.cls & 3u), (int)((g.cls >> 2) & 3u), mref(g.moff), mref(g.moff + 1)};
                    Coef fb{(int)((g.cls >> 4) & 3u), (int)((g.cls >> 6) & 3u), mref(g.moff + 2), mref(g.moff + 3)};
                    if (uni) {
                        o << "        const double " << un << "ar = su[" << 4 * g.uacc << "], " << un << "ai = su[" << 4 * g.uacc + 1
                          << "], " << un << "br = su[" << 4 * g.uacc + 2 << "], " << un << "bi = su[" << 4 * g.uacc + 3 << "];\n";
                        fa = Coef{CLS_GEN, CLS_GEN, un + "ar", un + "ai"};
                        fb = Coef{CLS_GEN, CLS_GEN, un + "br", un + "bi"};
                    }
                    else if (tz || g.zo || !cond.empty()) {
                        std::string pz = "false";
                        if (tz || g.zo) {
                            pz = "(((";
                            if (tz) pz += "__popc(ib & " + hex32(tz) + ")";
                            if (tz && g.zo) pz += " + ";
                            if (g.zo) pz += "__popcll(obase & " + hex64(g.zo) + ")";
                            pz += ") & 1) != 0)";
                        }
                        o << "        const bool " << un << "p = " << pz << ";\n";
                        const std::string c1 = cond.empty() ? "" : "!(" + cond + ") ? 1.0 : ";
                        const std::string c0 = cond.empty() ? "" : "!(" + cond + ") ? 0.0 : ";
                        o << "        const double " << un << "ar = " << c1 << "(" << un << "p ? " << mref(g.moff + 2) << " : " << mref(g.moff)
                          << "), " << un << "ai = " << c0 << "(" << un << "p ? " << mref(g.moff + 3) << " : " << mref(g.moff + 1) << ");\n";
                        o << "        const double " << un << "br = " << c1 << "(" << un << "p ? " << mref(g.moff) << " : " << mref(g.moff + 2)
                          << "), " << un << "bi = " << c0 << "(" << un << "p ? " << mref(g.moff + 1) << " : " << mref(g.moff + 3) << ");\n";
                        fa = Coef{CLS_GEN, CLS_GEN, un + "ar", un + "ai"};
                        fb = Coef{CLS_GEN, CLS_GEN, un + "br", un + "bi"};
                    }
                    Acc *A = nullptr;
                    for (auto &a : accs)
                        if (a.type == 1 && a.rcm == rcm && a.rcv == rcv && a.rz == rz) A = &a;
                    if (!A) {
                        accs.push_back(Acc{1, rcm, rcv, rz, fa, fb, (int)accs.size()});
                        continue;
                    }
                    if (A->a.cr != CLS_GEN || A->a.ci != CLS_GEN) std::swap(A->a, fa);
                    if (A->b.cr != CLS_GEN || A->b.ci != CLS_GEN) std::swap(A->b, fb);
                    mul_into(A->a, fa, "ga" + std::to_string(A->id) + "_" + std::to_string(u));
                    if (rz) mul_into(A->b, fb, "gb" + std::to_string(A->id) + "_" + std::to_string(u));
                }
            }
            // every slot times the product of the accumulators that apply to it (products are shared)
            std::vector<std::pair<std::vector<int>, Coef>> memo; // factor list (acc id * 2 + odd) -> value
            auto factor_of = [&](int code) -> Coef { return (code & 1) ? accs[code >> 1].b : accs[code >> 1].a; };
            for (uint32_t k = 0; k < nslots; k++) {
                std::vector<int> fl;
                for (auto &a : accs)
                    if ((pat[k] & a.rcm) == a.rcv)
                        fl.push_back(a.id * 2 + (a.type == 1 ? (__builtin_popcount(pat[k] & a.rz) & 1) : 0));
                if (fl.empty()) continue;
                Coef F = factor_of(fl[0]);
                std::vector<int> pre{fl[0]};
                for (size_t q = 1; q < fl.size(); q++) {
                    pre.push_back(fl[q]);
                    const Coef *hit = nullptr;
                    for (auto &m : memo)
                        if (m.first == pre) hit = &m.second;
                    if (hit) {
                        F = *hit;
                        continue;
                    }
                    Coef f = factor_of(fl[q]);
                    if (F.cr != CLS_GEN || F.ci != CLS_GEN) std::swap(F, f);
                    std::string nm = "p";
                    for (int c_ : pre) nm += "_" + std::to_string(c_);
                    mul_into(F, f, nm);
                    memo.push_back({pre, F});
                }
                std::vector<Term> re, im;
                cmul_terms(re, im, F, "ar", "ai");
                o << "        { const double ar = v" << k << "r, ai = v" << k << "i; v" << k << "r = " << lin(re) << "; v" << k
                  << "i = " << lin(im) << "; }\n";
            }
            o << "      }\n";
        };
        for (size_t oi = 0; oi < R.ops.size(); oi++) {
            const JOp &g = R.ops[oi];
            if (P.fuse_diag && (g.kind == GateOp::PHASE || g.kind == GateOp::PARITY)) {
                size_t oj = oi;
                while (oj < R.ops.size() && (R.ops[oj].kind == GateOp::PHASE || R.ops[oj].kind == GateOp::PARITY)) oj++;
                if (oj - oi >= 2) {
                    emit_diag_run(oi, oj);
                    oi = oj - 1;
                    continue;
                }
            }
            // run-time condition: outside-the-tile controls (uniform per tile) and thread-bit controls
            std::string cond;
            if (g.co_m) cond = "((obase & " + hex64(g.co_m) + ") == " + hex64(g.co_v) + ")";
            const uint32_t tcm = g.cin_m & ~R.S, tcv = g.cin_v & ~R.S;
            if (tcm) cond += std::string(cond.empty() ? "" : " && ") + "((ib & " + hex32(tcm) + ") == " + hex32(tcv) + ")";
            const uint32_t rcm = g.cin_m & R.S, rcv = g.cin_v & R.S;
            auto slot_on = [&](uint32_t k) { return (pat[k] & rcm) == rcv; };
            auto coef = [&](int idx) -> Coef { // complex coefficient idx of the op's parameter doubles
                return Coef{(int)((g.cls >> (4 * idx)) & 3u), (int)((g.cls >> (4 * idx + 2)) & 3u),
                            mref(g.moff + 2 * idx), mref(g.moff + 2 * idx + 1)};
            };
            std::ostringstream b; // op body
            if (g.kind == GateOp::PAIR) {
                const Coef m00 = coef(0), m01 = coef(1), m10 = coef(2), m11 = coef(3);
                for (uint32_t k = 0; k < nslots; k++) {
                    if ((pat[k] & g.x) || !slot_on(k)) continue;
                    const uint32_t A = slot_of(pat[k] | g.a), B = slot_of(pat[k] | g.b);
                    std::vector<Term> xr, xi, yr, yi;
                    cmul_terms(xr, xi, m00, "ar", "ai");
                    cmul_terms(xr, xi, m01, "br", "bi");
                    cmul_terms(yr, yi, m10, "ar", "ai");
                    cmul_terms(yr, yi, m11, "br", "bi");
                    b << "        { const double ar = " << V(A, 'r') << ", ai = " << V(A, 'i') << ", br = " << V(B, 'r')
                      << ", bi = " << V(B, 'i') << ";\n";
                    b << "          " << V(A, 'r') << " = " << lin(xr) << "; " << V(A, 'i') << " = " << lin(xi) << ";\n";
                    b << "          " << V(B, 'r') << " = " << lin(yr) << "; " << V(B, 'i') << " = " << lin(yi) << "; }\n";
                }
            }
            else if (g.kind == GateOp::PHASE) {
                const Coef e = coef(0);
                for (uint32_t k = 0; k < nslots; k++) {
                    if (!slot_on(k)) continue;
                    std::vector<Term> re, im;
                    cmul_terms(re, im, e, "ar", "ai");
                    b << "        { const double ar = " << V(k, 'r') << ", ai = " << V(k, 'i') << "; " << V(k, 'r') << " = "
                      << lin(re) << "; " << V(k, 'i') << " = " << lin(im) << "; }\n";
                }
            }
            else if (g.kind == GateOp::PARITY) {
                const uint32_t tz = g.zin & ~R.S, rz = g.zin & R.S;
                Coef e[2] = {coef(0), coef(1)};
                if (tz || g.zo) {
                    b << "        const bool pz_ = ((";
                    if (tz) b << "__popc(ib & " << hex32(tz) << ")";
                    if (tz && g.zo) b << " + ";
                    if (g.zo) b << "__popcll(obase & " << hex64(g.zo) << ")";
                    b << ") & 1) != 0;\n";
                    b << "        const double ear = pz_ ? " << mref(g.moff + 2) << " : " << mref(g.moff) << ", eai = pz_ ? "
                      << mref(g.moff + 3) << " : " << mref(g.moff + 1) << ", ebr = pz_ ? " << mref(g.moff) << " : "
                      << mref(g.moff + 2) << ", ebi = pz_ ? " << mref(g.moff + 1) << " : " << mref(g.moff + 3) << ";\n";
                    e[0] = Coef{CLS_GEN, CLS_GEN, "ear", "eai"};
                    e[1] = Coef{CLS_GEN, CLS_GEN, "ebr", "ebi"};
                }
                for (uint32_t k = 0; k < nslots; k++) {
                    if (!slot_on(k)) continue;
                    const Coef &ek = e[__builtin_popcount(pat[k] & rz) & 1];
                    std::vector<Term> re, im;
                    cmul_terms(re, im, ek, "ar", "ai");
                    b << "        { const double ar = " << V(k, 'r') << ", ai = " << V(k, 'i') << "; " << V(k, 'r') << " = "
                      << lin(re) << "; " << V(k, 'i') << " = " << lin(im) << "; }\n";
                }
            }
            else if (g.kind == GateOp::DENSE) { // 2^k x 2^k matrix on k register bits
                const int k = (int)g.tg.size(), dim = 1 << k;
                auto spread = [&](int r) {
                    uint32_t m = 0;
                    for (int i = 0; i < k; i++)
                        if ((r >> (k - 1 - i)) & 1) m |= 1u << g.tg[i];
                    return m;
                };
                for (uint32_t s0 = 0; s0 < nslots; s0++) {
                    if ((pat[s0] & g.x) || !slot_on(s0)) continue;
                    b << "        {\n";
                    for (int cc = 0; cc < dim; cc++) {
                        const uint32_t sl = slot_of(pat[s0] | spread(cc));
                        b << "          const double a" << cc << "r = " << V(sl, 'r') << ", a" << cc << "i = " << V(sl, 'i') << ";\n";
                    }
                    for (int r = 0; r < dim; r++) {
                        std::vector<Term> re, im;
                        for (int cc = 0; cc < dim; cc++) {
                            const int mo = g.moff + 2 * (r * dim + cc);
                            cmul_terms(re, im, Coef{CLS_GEN, CLS_GEN, mref(mo), mref(mo + 1)}, "a" + std::to_string(cc) + "r",
                                       "a" + std::to_string(cc) + "i");
                        }
                        const uint32_t sl = slot_of(pat[s0] | spread(r));
                        b << "          " << V(sl, 'r') << " = " << lin(re) << ";\n          " << V(sl, 'i') << " = " << lin(im) << ";\n";
                    }
                    b << "        }\n";
                }
            }
            else { // PAULIROT with all non-diagonal bits in registers
                const uint32_t tz = g.zin & ~R.S, rz = g.zin & R.S;
                const uint32_t top = 1u << (31 - __builtin_clz(g.x));
                std::string sn = mref(g.moff + 1);
                if (tz || g.zo) {
                    b << "        const bool pz_ = ((";
                    if (tz) b << "__popc(ib & " << hex32(tz) << ")";
                    if (tz && g.zo) b << " + ";
                    if (g.zo) b << "__popcll(obase & " << hex64(g.zo) << ")";
                    b << ") & 1) != 0;\n";
                    b << "        const double sn_ = pz_ ? -" << sn << " : " << sn << ";\n";
                    sn = "sn_";
                }
                const std::string cs = mref(g.moff);
                // W(v, e) = -i * i^e * v, as (re, im) expressions with sign
                auto W = [&](const char *r, const char *i_, int e, std::string &wr, bool &nr_, std::string &wi, bool &ni_) {
                    switch ((e + 3) & 3) {
                    case 0: wr = r; nr_ = false; wi = i_; ni_ = false; break;
                    case 1: wr = i_; nr_ = true; wi = r; ni_ = false; break;
                    case 2: wr = r; nr_ = true; wi = i_; ni_ = true; break;
                    default: wr = i_; nr_ = false; wi = r; ni_ = true; break;
                    }
                };
                for (uint32_t k = 0; k < nslots; k++) {
                    if ((pat[k] & top) || !slot_on(k)) continue;
                    const uint32_t A = k, B = slot_of((pat[k] ^ g.x) & R.S);
                    const int eA = g.ny + 2 * (__builtin_popcount(pat[A] & rz) & 1);
                    const int eB = g.ny + 2 * (__builtin_popcount(pat[B] & rz) & 1);
                    std::string wr, wi;
                    bool nr_, ni_;
                    b << "        { const double ar = " << V(A, 'r') << ", ai = " << V(A, 'i') << ", br = " << V(B, 'r')
                      << ", bi = " << V(B, 'i') << ";\n";
                    g_fp64_emitted += 8.0;
                    W("br", "bi", eB, wr, nr_, wi, ni_); // new_A = cs * A + sn * W(B, eB)
                    b << "          " << V(A, 'r') << " = fma(" << (nr_ ? "-" : "") << sn << ", " << wr << ", " << cs << " * ar); "
                      << V(A, 'i') << " = fma(" << (ni_ ? "-" : "") << sn << ", " << wi << ", " << cs << " * ai);\n";
                    W("ar", "ai", eA, wr, nr_, wi, ni_);
                    b << "          " << V(B, 'r') << " = fma(" << (nr_ ? "-" : "") << sn << ", " << wr << ", " << cs << " * br); "
                      << V(B, 'i') << " = fma(" << (ni_ ? "-" : "") << sn << ", " << wi << ", " << cs << " * bi); }\n";
                }
            }
            const std::string body = b.str();
            if (body.empty()) continue;
            if (!cond.empty()) o << "      if (" << cond << ") {\n" << body << "      }\n";
            else o << "      {\n" << body << "      }\n";
        }
        // Last round: the 2^nr amplitudes go from registers straight to global memory (16 STS + 16 LDS + one
        // barrier less per thread and tile: the light passes stall on the shared-memory instruction queue,
        // profiles/r2_v1_pass_summary.txt).  Lane bits sit on the lowest non-register tile positions, so a warp
        // still writes whole 32-byte sectors as long as tile position 0 is not a register bit.
        const bool last = rix == (int)P.rounds.size();
        direct_store = last && P.direct_store && !(R.S & 1u);
        if (direct_store) {
            o << "      const unsigned long long gib = 0ull";
            for (int k = 0; k < (int)qq.size();) {
                auto gpos = [&](int tp_) { return tp_ < c ? tp_ : P.high[tp_ - c]; };
                int len = 1;
                while (k + len < (int)qq.size() && qq[k + len] == qq[k] + len && gpos(qq[k + len]) == gpos(qq[k]) + len) len++;
                const uint32_t m = ((1u << len) - 1u) << k;
                const int sh = gpos(qq[k]) - k;
                o << " | ((unsigned long long)(tid & " << hex32(m) << ")" << (sh > 0 ? " << " + std::to_string(sh) : (sh < 0 ? " >> " + std::to_string(-sh) : "")) << ")";
                k += len;
            }
            o << ";\n";
            o << "      cplx *__restrict__ gp = psi + (base | gib);\n";
            for (uint32_t k = 0; k < nslots; k++)
                o << "      gp[" << hex64(off_of(pat[k])) << "] = make_double2(v" << k << "r, v" << k << "i);\n";
            o << "    }\n";
        }
        else {
            for (uint32_t k = 0; k < nslots; k++)
                o << "      s[" << slot_expr("sb", pat[k]) << "] = make_double2(v" << k << "r, v" << k << "i);\n";
            o << "    }\n    __syncthreads();\n";
        }
    }
    // ---- store -------------------------------------------------------------------------------------
    if (!direct_store)
        for (uint32_t k = 0; k < niter; k++)
            if (P.memhint & 1) o << "    __stcs(&tp[" << hex64(off_of(k * TT)) << "], s[" << slot_expr("swtid", k * TT) << "]);\n";
            else o << "    tp[" << hex64(off_of(k * TT)) << "] = s[" << slot_expr("swtid", k * TT) << "];\n";
    o << "    __syncthreads();\n"; // every thread is done with this tile's buffer before it is refilled
    if (P.dbuf) o << "    cur ^= 1u;\n";
    o << "  }\n}\n";

    K.name = name;
    K.threads = (int)TT;
    K.smem_bytes = (int)((P.dbuf ? 2 : 1) * 16 * nelem) + (P.uaccs.empty() ? 0 : 8 * (8 * P.nu + 4 * (int)P.uaccs.size()));
    K.param_doubles = 1 + std::max(1, P.nm);
    K.fp64_per_thread = g_fp64_emitted;
    return o.str();
}

} // namespace

// Returns false when the pass must stay on the interpreted kernel (JIT off or unavailable, kernel not
// compiled yet, an op that does not fit a register round, tile too small).
bool Simulator::launch_pass_jit(double *vec, const std::vector<const GateOp *> &ops, const std::vector<int> &high, int c)
{
    if (jit_mode_ == 0) return false;
    const int t = c + (int)high.size();
    // 'always' (tests) specialises every pass, however small; 'auto' only where it can pay off
    const bool always = jit_mode_ == 2;
    const int nr = std::min(opt_.jit_regbits, always ? t - 1 : t - 5);
    if (nr < (always ? 1 : 3) || t > 13 || (!always && nl_ < opt_.jit_min_qubits)) return false;
    jit::Engine &E = jit::Engine::get();
    if (dry_) B200_FAIL_IF(!E.nvrtc_available(), "dry run needs NVRTC");
    else if (!E.available()) {
        B200_FAIL_IF(jit_mode_ == 2, "jit='always' but NVRTC / the driver API could not be loaded");
        return false;
    }
    int loc[64];
    for (int p = 0; p < 64; p++) loc[p] = p < c ? p : -1;
    for (size_t h = 0; h < high.size(); h++) loc[high[h]] = c + (int)h;
    auto split = [&](uint64_t mask, uint32_t &inside, uint64_t &outside) {
        inside = 0;
        outside = 0;
        while (mask) {
            const int p = __builtin_ctzll(mask);
            mask &= mask - 1;
            if (loc[p] >= 0) inside |= 1u << loc[p];
            else outside |= 1ull << p;
        }
    };

    JPlan P;
    P.t = t;
    P.c = c;
    P.nr = nr;
    P.nl = (int)nl_;
    P.high = high;
    P.rank_base = rank_base_;
    P.fuse_diag = opt_.jit_fuse_diag;
    {
        const uint64_t ntiles = 1ull << (nl_ - t);
        const int per_sm = std::max(1, (227 * 1024) / (2 * 16 * (1 << t) + 1024));
        P.dbuf = opt_.jit_dbuf == 2 || // 2 = forced (tests)
                 (opt_.jit_dbuf && ntiles >= 2ull * (uint64_t)E.sm_count(cuda_device_) * (uint64_t)per_sm);
    }
    std::vector<double> mvals;
    std::vector<const GateOp *> rem(ops.begin(), ops.end()), deferred, taken;
    while (!rem.empty()) {
        uint64_t blockedX = 0, blockedZ = 0;
        uint32_t S = 0;
        deferred.clear();
        taken.clear();
        int ctrl_votes[32] = {0};
        for (const GateOp *g : rem) {
            const uint64_t xs = g->xmask, zs = g->zsupport();
            bool fits = false;
            if (!(xs & (blockedX | blockedZ)) && !(zs & blockedX)) {
                if (g->kind == GateOp::DENSE && ((int)g->targets.size() > nr || !g->matrix)) return false;
                uint32_t xin;
                uint64_t xout;
                split(xs, xin, xout);
                if (xout) return false;
                if (__builtin_popcount(xin) > nr) return false;
                if (__builtin_popcount(S | xin) <= nr) {
                    S |= xin;
                    fits = true;
                }
            }
            if (fits) {
                taken.push_back(g);
                if (g->kind != GateOp::PHASE && g->kind != GateOp::PARITY) { // controls of non-diagonal ops
                    uint32_t cin;
                    uint64_t cout_;
                    split(g->cmask, cin, cout_);
                    for (int b = 0; b < t; b++)
                        if ((cin >> b) & 1u) ctrl_votes[b]++;
                }
            }
            else {
                deferred.push_back(g);
                blockedX |= xs;
                blockedZ |= zs;
            }
        }
        // Free register slots go to the control bits of the round's non-diagonal ops first: a control on a
        // register bit is a compile-time slot selection (CNOT/Toffoli become pure renames), a control on a
        // thread bit is a run-time condition whose permutation costs ~6 register moves per amplitude.
        while (__builtin_popcount(S) < nr) {
            int best = -1;
            for (int b = 0; b < t; b++)
                if (!((S >> b) & 1u) && ctrl_votes[b] > 0 && (best < 0 || ctrl_votes[b] > ctrl_votes[best])) best = b;
            if (best < 0) break;
            S |= 1u << best;
        }
        for (int b = t - 1; b >= 0 && __builtin_popcount(S) < nr; b--) // pad with the highest free tile bits
            if (!(S & (1u << b))) S |= 1u << b;
        JRound R;
        R.S = S;
        for (const GateOp *gp : taken) {
            const GateOp &g = *gp;
            JOp o;
            o.kind = (int)g.kind;
            uint64_t dummy;
            split(g.xmask, o.x, dummy);
            split(g.a, o.a, dummy);
            split(g.b, o.b, dummy);
            split(g.cmask, o.cin_m, o.co_m);
            split(g.cval, o.cin_v, o.co_v);
            if (g.kind == GateOp::PARITY || g.kind == GateOp::PAULIROT) split(g.zmask, o.zin, o.zo);
            o.ny = g.ny & 3;
            switch (g.kind) {
            case GateOp::DENSE: {
                const size_t dim = size_t(1) << g.targets.size();
                for (int tb : g.targets) o.tg.push_back(loc[tb]);
                o.nm = (int)(2 * dim * dim);
                o.dvals.resize((size_t)o.nm);
                for (size_t q = 0; q < dim * dim; q++) {
                    o.dvals[2 * q] = (*g.matrix)[q].real();
                    o.dvals[2 * q + 1] = (*g.matrix)[q].imag();
                }
                break;
            }
            case GateOp::PAIR:
                o.nm = 8;
                for (int q = 0; q < 4; q++) {
                    o.vals[2 * q] = g.m[q].real();
                    o.vals[2 * q + 1] = g.m[q].imag();
                }
                break;
            case GateOp::PHASE:
                o.nm = 2;
                o.vals[0] = g.m[0].real();
                o.vals[1] = g.m[0].imag();
                break;
            case GateOp::PARITY:
                o.nm = 4;
                o.vals[0] = g.m[0].real(); o.vals[1] = g.m[0].imag();
                o.vals[2] = g.m[1].real(); o.vals[3] = g.m[1].imag();
                break;
            default: // PAULIROT: (cos, sin)
                o.nm = 2;
                o.vals[0] = g.m[0].real();
                o.vals[1] = g.m[0].imag();
                break;
            }
            o.literal = g.kind == GateOp::PAIR || g.kind == GateOp::PHASE ||
                        (g.kind == GateOp::PARITY && !(o.zin & ~S) && !o.zo);
            R.ops.push_back(o);
        }
        P.rounds.push_back(std::move(R));
        rem.swap(deferred);
    }
    if (opt_.jit_factor) factor_scales(P);
    P.uniform = opt_.jit_uniform;
    P.direct_store = opt_.jit_direct_store;
    P.memhint = opt_.jit_memhint;
    analyze_uniform(P);
    for (JRound &R : P.rounds)
        for (JOp &o : R.ops) {
            o.cls = 0;
            if (o.uacc >= 0) continue; // its values travel in the uniform table, below
            o.moff = (int)mvals.size();
            if (o.kind == (int)GateOp::DENSE) mvals.insert(mvals.end(), o.dvals.begin(), o.dvals.end());
            else {
                for (int q = 0; q < o.nm; q++) o.cls |= (uint32_t)(o.literal ? cls_of(o.vals[q]) : CLS_GEN) << (2 * q);
                for (int q = 0; q < o.nm; q++) mvals.push_back(o.vals[q]);
            }
        }
    // uniform table: entries of one accumulator are contiguous; 8 doubles per op =
    // { predicate mask, predicate value, z mask (bit patterns), 0, e0.re, e0.im, e1.re, e1.im }
    P.uoff = (int)mvals.size();
    if (!P.uaccs.empty()) {
        mvals.resize((size_t)P.uoff + 8 * (size_t)P.nu, 0.0);
        std::vector<int> fill(P.uaccs.size(), 0);
        for (JRound &R : P.rounds)
            for (JOp &o : R.ops) {
                if (o.uacc < 0) continue;
                const int e = P.uaccs[(size_t)o.uacc].first + fill[(size_t)o.uacc]++;
                double *dst = &mvals[(size_t)P.uoff + 8 * (size_t)e];
                const uint64_t zm = o.kind == (int)GateOp::PARITY ? o.zo : 0ull;
                std::memcpy(dst + 0, &o.co_m, 8);
                std::memcpy(dst + 1, &o.co_v, 8);
                std::memcpy(dst + 2, &zm, 8);
                dst[4] = o.vals[0];
                dst[5] = o.vals[1];
                dst[6] = o.kind == (int)GateOp::PARITY ? o.vals[2] : o.vals[0];
                dst[7] = o.kind == (int)GateOp::PARITY ? o.vals[3] : o.vals[1];
                o.moff = P.uoff + 8 * e;
            }
    }
    P.nm = (int)mvals.size();
    if (P.nm * 8 + 24 > 32000) return false; // kernel parameter block limit

    // structural key
    std::string key;
    key.reserve(64 + 64 * ops.size());
    put(key, (int)10); // generator version
    put(key, P.fuse_diag);
    put(key, P.uniform);
    put(key, P.direct_store);
    put(key, P.memhint);
    put(key, P.t); put(key, P.c); put(key, P.nr); put(key, P.nl); put(key, P.dbuf);
    for (int h : P.high) put(key, (uint8_t)h);
    for (const JRound &R : P.rounds) {
        put(key, (uint32_t)0xffffffffu);
        put(key, R.S);
        for (const JOp &o : R.ops) {
            put(key, (uint8_t)o.kind); put(key, (uint8_t)o.ny);
            put(key, o.x); put(key, o.a); put(key, o.b); put(key, o.cin_m); put(key, o.cin_v);
            put(key, o.co_m); put(key, o.co_v); put(key, o.zin); put(key, o.zo); put(key, o.cls);
            put(key, (int16_t)o.uacc);
            for (int tb : o.tg) put(key, (uint8_t)tb);
        }
    }
    auto gen = [&](jit::Kernel &k) {
        const uint64_t h = std::hash<std::string>{}(key);
        char name[64];
        snprintf(name, sizeof name, "k_pass_%016llx", (unsigned long long)h);
        return generate(P, k, name);
    };
    if (dry_) { // development: compile to a cubin once per structure, launch nothing
        static std::map<std::string, double> seen;
        auto it = seen.find(key);
        if (it == seen.end()) {
            jit::Kernel k;
            std::string src;
            B200_FAIL_IF(!E.compile_offline(key, gen, k, src), "dry run: NVRTC failed: " + k.error);
            it = seen.emplace(key, k.fp64_per_thread * (double)k.threads).first;
            if (std::getenv("B200SV_JIT_VERBOSE"))
                fprintf(stderr, "[b200sv dry] %s: t=%d c=%d rounds=%zu ops=%zu fp64/thread=%.0f %.0f ms\n", k.name.c_str(), t, c,
                        P.rounds.size(), ops.size(), k.fp64_per_thread, k.compile_ms);
        }
        stats_.tile_passes++;
        stats_.jit_passes++;
        stats_.kernel_ops += (int64_t)ops.size();
        stats_.rounds += (int64_t)P.rounds.size();
        stats_.fp64_ops += it->second * (double)(1ull << (nl_ - t));
        return true;
    }
    // a state of >= 1 GiB pays ~20 ms per interpreted pass: compile at the first sighting (in the background)
    const int threshold = jit_mode_ == 2 ? 1 : (nl_ >= 26 ? 1 : opt_.jit_threshold);
    const jit::Kernel *K = E.lookup(cuda_device_, key, threshold, jit_mode_ == 2, gen);
    if (!K) {
        B200_FAIL_IF(jit_mode_ == 2, "jit='always': specialised pass failed to compile (set B200SV_JIT_VERBOSE=1)");
        return false;
    }
    const uint64_t ntiles = 1ull << (nl_ - t);
    uint64_t grid = (uint64_t)E.sm_count(cuda_device_) * (uint64_t)K->ctas_per_sm;
    if (grid > ntiles) grid = ntiles;
    if (mvals.empty()) mvals.push_back(0.0);
    std::vector<double> blk(1 + mvals.size());
    std::memcpy(&blk[0], &rank_base_, sizeof(double)); // Params::rank_base
    std::memcpy(&blk[1], mvals.data(), mvals.size() * sizeof(double));
    // $B200SV_PASS_LOG=1 (analysis): one stderr line per specialised pass with its CUDA-event time
    static const bool pass_log = std::getenv("B200SV_PASS_LOG") != nullptr;
    if (pass_log) {
        if (!ev2_) {
            cudaEvent_t a, b;
            cudaEventCreate(&a);
            cudaEventCreate(&b);
            ev2_ = a;
            ev3_ = b;
        }
        cudaEventRecord((cudaEvent_t)ev2_, (cudaStream_t)stream_);
    }
    const int rc = E.launch(*K, (unsigned)grid, vec, blk.data(), stream_);
    B200_FAIL_IF(rc != 0, "cuLaunchKernel of a specialised pass failed with driver error " + std::to_string(rc));
    if (pass_log) {
        cudaEventRecord((cudaEvent_t)ev3_, (cudaStream_t)stream_);
        cudaEventSynchronize((cudaEvent_t)ev3_);
        float ms = 0;
        cudaEventElapsedTime(&ms, (cudaEvent_t)ev2_, (cudaEvent_t)ev3_);
        int nuni = 0;
        for (const JRound &R : P.rounds)
            for (const JOp &o : R.ops) nuni += o.uacc >= 0;
        fprintf(stderr, "[b200sv pass] %s t=%d c=%d rounds=%zu ops=%zu uniform_ops=%d uaccs=%zu fp64/thread=%.0f regs=%d ctas/sm=%d %.3f ms %.0f GB/s\n",
                K->name.c_str(), t, c, P.rounds.size(), ops.size(), nuni, P.uaccs.size(), K->fp64_per_thread, K->num_regs,
                K->ctas_per_sm, ms, 32.0 * std::ldexp(1.0, (int)nl_) / (ms * 1e-3) / 1e9);
    }
    stats_.tile_passes++;
    stats_.jit_passes++;
    stats_.kernel_ops += (int64_t)ops.size();
    stats_.rounds += (int64_t)P.rounds.size();
    stats_.fp64_ops += K->fp64_per_thread * (double)K->threads * (double)ntiles;
    stats_.h2d_bytes += (double)(8 + blk.size() * sizeof(double)); // the kernel parameter block
    return true;
}

b200sv_stats Simulator::stats() const
{
    const_cast<Simulator *>(this)->collect_timing(); // resolves the CUDA-event pairs of the flushes so far (waits for them)
    b200sv_stats s = stats_;
    if (jit_mode_ != 0) {
        const jit::Counters c = jit::Engine::get().counters();
        s.jit_compiled = c.compiled;
        s.jit_pending = c.pending;
        s.jit_compile_ms = c.compile_ms;
        s.jit_disk_hits = c.disk_hits;
    }
    return s;
}

void jit_quiesce() { jit::Engine::get().quiesce(); }

void Simulator::jit_wait()
{
    if (jit_mode_ != 0) jit::Engine::get().wait_all();
}

} // namespace b200
