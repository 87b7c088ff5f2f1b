// ORACLE — TEST INFRASTRUCTURE ONLY (see sv_oracle.hpp header).  Flat C ABI over orc::Oracle so the
// Python tests and bench.py's cpu_baseline leg can drive the CPU restatement through ctypes.
#include "sv_oracle.hpp"

#include <cstring>

#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

static thread_local std::string g_err;

#define ORC_TRY(...)                                                                               \
    try {                                                                                          \
        __VA_ARGS__;                                                                               \
        return 0;                                                                                  \
    }                                                                                              \
    catch (const std::exception &e) {                                                              \
        g_err = e.what();                                                                          \
        return -1;                                                                                 \
    }

static std::vector<size_t> wv(const int64_t *w, int n)
{
    std::vector<size_t> r(n);
    for (int i = 0; i < n; i++) {
        ORC_FAIL_IF(w[i] < 0, "Invalid wire");
        r[i] = size_t(w[i]);
    }
    return r;
}
static std::vector<bool> bv(const uint8_t *v, int n)
{
    std::vector<bool> r(n);
    for (int i = 0; i < n; i++) r[i] = v[i] != 0;
    return r;
}
static CVec cv(const double *ri, size_t n)
{
    CVec r(n);
    for (size_t i = 0; i < n; i++) r[i] = cplx(ri[2 * i], ri[2 * i + 1]);
    return r;
}

extern "C" {

const char *orc_last_error() { return g_err.c_str(); }
int orc_num_threads()
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void orc_set_num_threads(int t)
{
#ifdef _OPENMP
    omp_set_num_threads(t);
#else
    (void)t;
#endif
}

void *orc_create(int n)
{
    try {
        return new Oracle(size_t(n));
    }
    catch (const std::exception &e) {
        g_err = e.what();
        return nullptr;
    }
}
void orc_destroy(void *h) { delete static_cast<Oracle *>(h); }
int orc_reset(void *h, int n) { ORC_TRY(static_cast<Oracle *>(h)->reset(size_t(n))) }
int orc_allocate_qubit(void *h) { ORC_TRY(static_cast<Oracle *>(h)->allocate_qubit()) }
int orc_release_qubit(void *h, int64_t wire) { ORC_TRY(static_cast<Oracle *>(h)->release_qubit(size_t(wire))) }
int orc_num_qubits(void *h) { return int(static_cast<Oracle *>(h)->n); }
void orc_seed(void *h, uint32_t s) { static_cast<Oracle *>(h)->seed(s); }
void orc_unseed(void *h) { static_cast<Oracle *>(h)->unseed(); }
void orc_set_shots(void *h, int64_t shots) { static_cast<Oracle *>(h)->shots = size_t(shots < 0 ? 0 : shots); }
void orc_set_seed_every_measurement(void *h, int on)
{
    static_cast<Oracle *>(h)->seed_every_measurement = on != 0;
}
// next raw 32-bit output of the runtime generator (for stream-position checks)
uint32_t orc_gen_peek(void *h)
{
    auto *o = static_cast<Oracle *>(h);
    if (!o->gen) return 0;
    std::mt19937 copy = *o->gen;
    return uint32_t(copy());
}

int orc_named_op(void *h, const char *name, const double *params, int np, const int64_t *wires,
                 int nw, int inverse, const int64_t *cw, const uint8_t *cvals, int nc,
                 const char *pauli_word)
{
    ORC_TRY(static_cast<Oracle *>(h)->named_op(name, std::vector<double>(params, params + np),
                                               wv(wires, nw), inverse != 0, wv(cw, nc),
                                               bv(cvals, nc), pauli_word ? pauli_word : ""))
}
int orc_matrix_op(void *h, const double *m_ri, const int64_t *wires, int nw, int inverse,
                  const int64_t *cw, const uint8_t *cvals, int nc)
{
    ORC_TRY(static_cast<Oracle *>(h)->matrix_op(cv(m_ri, (size_t(1) << nw) * (size_t(1) << nw)),
                                                wv(wires, nw), inverse != 0, wv(cw, nc),
                                                bv(cvals, nc)))
}
int orc_set_state(void *h, const double *st_ri, const int64_t *wires, int nw)
{
    ORC_TRY(static_cast<Oracle *>(h)->set_state(cv(st_ri, size_t(1) << nw), wv(wires, nw)))
}
int orc_set_basis_state(void *h, const int8_t *bits, const int64_t *wires, int nw)
{
    ORC_TRY(static_cast<Oracle *>(h)->set_basis_state(std::vector<int8_t>(bits, bits + nw),
                                                      wv(wires, nw)))
}
int orc_state(void *h, double *out_ri)
{
    auto *o = static_cast<Oracle *>(h);
    std::memcpy(out_ri, o->psi.data(), o->psi.size() * sizeof(cplx));
    return 0;
}

int64_t orc_obs_named(void *h, int id, const int64_t *wires, int nw)
{
    try {
        return static_cast<Oracle *>(h)->obs_named(id, wv(wires, nw));
    }
    catch (const std::exception &e) {
        g_err = e.what();
        return -1;
    }
}
int64_t orc_obs_hermitian(void *h, const double *m_ri, const int64_t *wires, int nw)
{
    try {
        return static_cast<Oracle *>(h)->obs_hermitian(
            cv(m_ri, (size_t(1) << nw) * (size_t(1) << nw)), wv(wires, nw));
    }
    catch (const std::exception &e) {
        g_err = e.what();
        return -1;
    }
}
int64_t orc_obs_tensor(void *h, const int64_t *keys, int nk)
{
    try {
        return static_cast<Oracle *>(h)->obs_tensor(std::vector<int64_t>(keys, keys + nk));
    }
    catch (const std::exception &e) {
        g_err = e.what();
        return -1;
    }
}
int64_t orc_obs_hamiltonian(void *h, const double *coeffs, const int64_t *keys, int nk)
{
    try {
        return static_cast<Oracle *>(h)->obs_hamiltonian(std::vector<double>(coeffs, coeffs + nk),
                                                         std::vector<int64_t>(keys, keys + nk));
    }
    catch (const std::exception &e) {
        g_err = e.what();
        return -1;
    }
}
int orc_expval(void *h, int64_t key, double *out) { ORC_TRY(*out = static_cast<Oracle *>(h)->expval(key)) }
int orc_var(void *h, int64_t key, double *out) { ORC_TRY(*out = static_cast<Oracle *>(h)->var(key)) }
int orc_probs(void *h, const int64_t *wires, int nw, double *out)
{
    ORC_TRY(auto p = static_cast<Oracle *>(h)->probs(wv(wires, nw));
            std::memcpy(out, p.data(), p.size() * sizeof(double)))
}
int orc_set_mcmc(void *h, int on, int64_t num_burnin, const char *kernel_name)
{
    ORC_TRY(auto *o = static_cast<Oracle *>(h); o->mcmc = on != 0; o->num_burnin = size_t(num_burnin);
            o->kernel_name = kernel_name ? kernel_name : "Local")
}
int orc_sample(void *h, const int64_t *wires, int nw, int64_t nshots, double *out)
{
    ORC_TRY(auto s = static_cast<Oracle *>(h)->sample(wv(wires, nw), size_t(nshots));
            std::memcpy(out, s.data(), s.size() * sizeof(double)))
}
int orc_counts(void *h, const int64_t *wires, int nw, int64_t nshots, double *eig, int64_t *cnt)
{
    ORC_TRY(std::vector<double> e; std::vector<int64_t> c;
            static_cast<Oracle *>(h)->counts(wv(wires, nw), size_t(nshots), e, c);
            std::memcpy(eig, e.data(), e.size() * sizeof(double));
            std::memcpy(cnt, c.data(), c.size() * sizeof(int64_t)))
}
int orc_measure(void *h, int64_t wire, int postselect, int *out)
{
    ORC_TRY(*out = static_cast<Oracle *>(h)->measure(size_t(wire), postselect) ? 1 : 0)
}
int orc_pauli_measure(void *h, const char *word, const int64_t *wires, int nw, int postselect, int *out)
{
    ORC_TRY(std::vector<size_t> w; for (int i = 0; i < nw; i++) w.push_back(size_t(wires[i]));
            *out = static_cast<Oracle *>(h)->pauli_measure(std::string(word), w, postselect) ? 1 : 0)
}
int orc_start_tape(void *h) { ORC_TRY(static_cast<Oracle *>(h)->start_tape()) }
int orc_stop_tape(void *h) { ORC_TRY(static_cast<Oracle *>(h)->stop_tape()) }
int orc_tape_num_params(void *h)
{
    size_t P = 0;
    for (auto &op : static_cast<Oracle *>(h)->tape) P += op.params.size();
    return int(P);
}
int orc_tape_num_obs(void *h) { return int(static_cast<Oracle *>(h)->tape_obs.size()); }
// out: n_obs x n_train row-major; train may be null (ntrain = 0 => all params)
int orc_gradient(void *h, const int64_t *train, int ntrain, double *out)
{
    ORC_TRY(auto j = static_cast<Oracle *>(h)->gradient(wv(train, ntrain));
            std::memcpy(out, j.data(), j.size() * sizeof(double)))
}

} // extern "C"
