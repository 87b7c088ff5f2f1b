// plugin.cpp — Catalyst device plugin: the C++ vtable adapter between Catalyst's runtime and the
// C ABI of include/b200sv.h.
//
// Catalyst's runtime dlopen()s this library and calls `B200StateVectorFactory(kwargs)`
// (reference: runtime/include/QuantumDevice.hpp:30-33, runtime/lib/capi/ExecutionContext.hpp:
// 254-267); every `__catalyst__qis__*` / `__catalyst__rt__*` call of a compiled program then lands
// on one of the virtual methods below (caller: runtime/lib/capi/RuntimeCAPI.cpp:241-1106).
// The adapter holds no numerics: it maps the runtime's logical qubit ids to device wires
// (reference helper QubitManager, runtime/lib/backend/common/QubitManager.hpp:34-132 — an
// interface-level header, included unchanged), unpacks DataView strides (runtime/include/
// DataView.hpp:124-134) and turns C-ABI error codes into RuntimeException (runtime/include/
// Exception.hpp:26-86).  It is compiled against the reference's headers where they are present
// (see catalyst_b200/build.py); nothing of the reference is copied into this repository.
#include <complex>
#include <cstdint>
#include <optional>
#include <random>
#include <string>
#include <vector>

#include "QuantumDevice.hpp"
#include "QubitManager.hpp"

#include "b200sv.h"

namespace Catalyst::Runtime::Devices {

#define B200_RT_CHECK(expr)                                                                        \
    do {                                                                                           \
        if ((expr) != 0) RT_FAIL(b200sv_last_error());                                             \
    } while (0)

class B200StateVector final : public Catalyst::Runtime::QuantumDevice {
  private:
    b200sv_device *dev_{nullptr};
    Catalyst::Runtime::QubitManager<QubitIdType, size_t> qubit_manager_{};
    std::mt19937 *gen_{nullptr};
    bool static_true_{true}, static_false_{false}; // Result storage that outlives Measure()

    static uint32_t next_u32(void *ctx) { return (uint32_t)(*static_cast<std::mt19937 *>(ctx))(); }

    auto dev_wires(const std::vector<QubitIdType> &wires) -> std::vector<int64_t>
    {
        RT_FAIL_IF(!qubit_manager_.isValidQubitId(wires), "Given wires do not refer to qubits");
        std::vector<int64_t> out;
        out.reserve(wires.size());
        for (auto w : wires) out.push_back((int64_t)qubit_manager_.getDeviceId(w));
        return out;
    }
    static auto ctrl_bytes(const std::vector<bool> &v) -> std::vector<uint8_t>
    {
        std::vector<uint8_t> out(v.size());
        for (size_t i = 0; i < v.size(); i++) out[i] = v[i] ? 1 : 0;
        return out;
    }
    template <typename T> static int64_t stride1(DataView<T, 1> &v)
    {
        return v.size() > 1 ? (int64_t)(&v(1) - &v(0)) : 1;
    }

  public:
    explicit B200StateVector(const std::string &kwargs = "{}")
    {
        dev_ = b200sv_create(kwargs.c_str());
        RT_FAIL_IF(dev_ == nullptr, b200sv_last_error());
    }
    ~B200StateVector() override { b200sv_destroy(dev_); }

    // ---- qubit management (QuantumDevice.hpp:88-144) -----------------------------------------
    auto AllocateQubits(size_t num_qubits) -> std::vector<QubitIdType> override
    {
        if (num_qubits == 0) return {};
        const size_t cur = (size_t)b200sv_num_qubits(dev_);
        if (cur == 0) {
            B200_RT_CHECK(b200sv_allocate_qubits(dev_, (int64_t)num_qubits));
        }
        else { // dynamic growth: new wires are appended in |0>
            for (size_t i = 0; i < num_qubits; i++) B200_RT_CHECK(b200sv_allocate_qubit(dev_));
        }
        return qubit_manager_.AllocateRange(cur, num_qubits);
    }
    auto AllocateQubit() -> QubitIdType override
    {
        const size_t cur = (size_t)b200sv_num_qubits(dev_);
        B200_RT_CHECK(b200sv_allocate_qubit(dev_));
        return qubit_manager_.Allocate(cur);
    }
    void ReleaseQubits(const std::vector<QubitIdType> &qubits) override
    {
        RT_FAIL_IF(!qubit_manager_.isValidQubitId(qubits), "Given wires do not refer to qubits");
        if (qubits.size() == (size_t)b200sv_num_qubits(dev_)) {
            // Releasing the full register returns the (pooled, reusable) object to a clean state
            // (runtime/lib/capi/ExecutionContext.hpp:355-361).
            B200_RT_CHECK(b200sv_release_qubits(dev_));
            qubit_manager_.ReleaseAll();
            return;
        }
        // a dynamically allocated sub-register (qml.allocate / deallocate): one wire at a time, each
        // looked up again because the earlier releases moved the later device wires down
        for (auto q : qubits) ReleaseQubit(q);
    }
    void ReleaseQubit(QubitIdType qubit) override
    {
        RT_FAIL_IF(!qubit_manager_.isValidQubitId(qubit), "Given wire does not refer to a qubit");
        if (b200sv_num_qubits(dev_) == 1) {
            B200_RT_CHECK(b200sv_release_qubits(dev_));
            qubit_manager_.ReleaseAll();
            return;
        }
        B200_RT_CHECK(b200sv_release_qubit(dev_, (int64_t)qubit_manager_.getDeviceId(qubit)));
        qubit_manager_.Release(qubit); // later device indices move down (QubitManager.hpp:124-126)
    }
    auto GetNumQubits() const -> size_t override { return (size_t)b200sv_num_qubits(dev_); }

    // ---- execution configuration (QuantumDevice.hpp:164-192) -----------------------------------
    void SetDeviceShots(size_t shots) override { B200_RT_CHECK(b200sv_set_shots(dev_, (int64_t)shots)); }
    auto GetDeviceShots() const -> size_t override { return (size_t)b200sv_get_shots(dev_); }
    void SetDevicePRNG(std::mt19937 *gen) override
    {
        gen_ = gen;
        B200_RT_CHECK(b200sv_set_prng(dev_, gen ? &next_u32 : nullptr, gen));
    }

    // ---- gates (QuantumDevice.hpp:215-284) -----------------------------------------------------
    void NamedOperation(const std::string &name, const std::vector<double> &params,
                        const std::vector<QubitIdType> &wires, bool inverse = false,
                        const std::vector<QubitIdType> &controlled_wires = {},
                        const std::vector<bool> &controlled_values = {},
                        const std::vector<std::string> &optional_params = {}) override
    {
        const auto w = dev_wires(wires), cw = dev_wires(controlled_wires);
        const auto cv = ctrl_bytes(controlled_values);
        RT_FAIL_IF(cw.size() != cv.size(), "Controlled wires/values size mismatch");
        const char *word = optional_params.empty() ? nullptr : optional_params[0].c_str();
        B200_RT_CHECK(b200sv_named_operation(dev_, name.c_str(), params.data(), (int64_t)params.size(),
                                             w.data(), (int64_t)w.size(), inverse ? 1 : 0, cw.data(),
                                             cv.data(), (int64_t)cw.size(), word));
    }
    void MatrixOperation(const std::vector<std::complex<double>> &matrix,
                         const std::vector<QubitIdType> &wires, bool inverse = false,
                         const std::vector<QubitIdType> &controlled_wires = {},
                         const std::vector<bool> &controlled_values = {}) override
    {
        const auto w = dev_wires(wires), cw = dev_wires(controlled_wires);
        const auto cv = ctrl_bytes(controlled_values);
        RT_FAIL_IF(cw.size() != cv.size(), "Controlled wires/values size mismatch");
        RT_FAIL_IF(w.empty() || w.size() > 10, "Invalid number of wires for MatrixOperation");
        const size_t dim = size_t(1) << w.size();
        RT_FAIL_IF(matrix.size() != dim * dim, "Invalid given matrix size");
        B200_RT_CHECK(b200sv_matrix_operation(dev_, reinterpret_cast<const double *>(matrix.data()),
                                              w.data(), (int64_t)w.size(), inverse ? 1 : 0, cw.data(),
                                              cv.data(), (int64_t)cw.size()));
    }
    void SetBasisState(DataView<int8_t, 1> &n, std::vector<QubitIdType> &wires) override
    {
        const auto w = dev_wires(wires);
        RT_FAIL_IF(n.size() != w.size(), "BasisState parameter and wires must be of equal length.");
        B200_RT_CHECK(b200sv_set_basis_state(dev_, n.size() ? &n(0) : nullptr, stride1(n), w.data(),
                                             (int64_t)w.size()));
    }
    void SetState(DataView<std::complex<double>, 1> &state, std::vector<QubitIdType> &wires) override
    {
        const auto w = dev_wires(wires);
        RT_FAIL_IF(state.size() != (size_t(1) << w.size()), "Invalid state size: must be 2^len(wires)");
        B200_RT_CHECK(b200sv_set_state(dev_, reinterpret_cast<const double *>(&state(0)),
                                       stride1(state), w.data(), (int64_t)w.size()));
    }

    // ---- observables (QuantumDevice.hpp:317-353) -------------------------------------------------
    auto Observable(ObsId id, const std::vector<std::complex<double>> &matrix,
                    const std::vector<QubitIdType> &wires) -> ObsIdType override
    {
        const auto w = dev_wires(wires);
        const int64_t key = b200sv_observable(dev_, (int)id, reinterpret_cast<const double *>(matrix.data()),
                                              (int64_t)matrix.size(), w.data(), (int64_t)w.size());
        RT_FAIL_IF(key < 0, b200sv_last_error());
        return (ObsIdType)key;
    }
    auto TensorObservable(const std::vector<ObsIdType> &obs) -> ObsIdType override
    {
        std::vector<int64_t> keys(obs.begin(), obs.end());
        const int64_t key = b200sv_tensor_observable(dev_, keys.data(), (int64_t)keys.size());
        RT_FAIL_IF(key < 0, b200sv_last_error());
        return (ObsIdType)key;
    }
    auto HamiltonianObservable(const std::vector<double> &coeffs, const std::vector<ObsIdType> &obs)
        -> ObsIdType override
    {
        RT_FAIL_IF(coeffs.size() != obs.size(), "Invalid size for the list of coefficients and observables");
        std::vector<int64_t> keys(obs.begin(), obs.end());
        const int64_t key =
            b200sv_hamiltonian_observable(dev_, coeffs.data(), keys.data(), (int64_t)keys.size());
        RT_FAIL_IF(key < 0, b200sv_last_error());
        return (ObsIdType)key;
    }

    // ---- measurements (QuantumDevice.hpp:234, 384-518) ---------------------------------------------
    auto Expval(ObsIdType obsKey) -> double override
    {
        double out = 0;
        B200_RT_CHECK(b200sv_expval(dev_, (int64_t)obsKey, &out));
        return out;
    }
    auto Var(ObsIdType obsKey) -> double override
    {
        double out = 0;
        B200_RT_CHECK(b200sv_var(dev_, (int64_t)obsKey, &out));
        return out;
    }
    void State(DataView<std::complex<double>, 1> &state) override
    {
        RT_FAIL_IF(state.size() != (size_t(1) << GetNumQubits()), "Invalid size for the pre-allocated state vector");
        B200_RT_CHECK(b200sv_state(dev_, reinterpret_cast<double *>(&state(0)), stride1(state)));
    }
    void Probs(DataView<double, 1> &probs) override
    {
        RT_FAIL_IF(probs.size() != (size_t(1) << GetNumQubits()), "Invalid size for the pre-allocated probabilities");
        B200_RT_CHECK(b200sv_probs(dev_, nullptr, 0, &probs(0), stride1(probs)));
    }
    void PartialProbs(DataView<double, 1> &probs, const std::vector<QubitIdType> &wires) override
    {
        const auto w = dev_wires(wires);
        RT_FAIL_IF(w.empty() || w.size() > GetNumQubits(), "Invalid number of wires");
        RT_FAIL_IF(probs.size() != (size_t(1) << w.size()), "Invalid size for the pre-allocated partial-probabilities");
        B200_RT_CHECK(b200sv_probs(dev_, w.data(), (int64_t)w.size(), &probs(0), stride1(probs)));
    }
    void sample_impl(DataView<double, 2> &samples, const std::vector<int64_t> &w)
    {
        const size_t shots = GetDeviceShots();
        const size_t k = w.empty() ? GetNumQubits() : w.size();
        RT_FAIL_IF(samples.size() != shots * k, "Invalid size for the pre-allocated samples");
        if (shots == 0 || k == 0) return;
        double *base = &samples(0, 0);
        const int64_t s0 = shots > 1 ? (int64_t)(&samples(1, 0) - base) : (int64_t)k;
        const int64_t s1 = k > 1 ? (int64_t)(&samples(0, 1) - base) : 1;
        B200_RT_CHECK(b200sv_sample(dev_, w.data(), (int64_t)w.size(), (int64_t)shots, base, s0, s1));
    }
    void Sample(DataView<double, 2> &samples) override { sample_impl(samples, {}); }
    void PartialSample(DataView<double, 2> &samples, const std::vector<QubitIdType> &wires) override
    {
        const auto w = dev_wires(wires);
        RT_FAIL_IF(w.empty() || w.size() > GetNumQubits(), "Invalid number of wires");
        sample_impl(samples, w);
    }
    void counts_impl(DataView<double, 1> &eigvals, DataView<int64_t, 1> &counts, const std::vector<int64_t> &w)
    {
        const size_t k = w.empty() ? GetNumQubits() : w.size();
        RT_FAIL_IF(eigvals.size() != (size_t(1) << k) || counts.size() != (size_t(1) << k),
                   "Invalid size for the pre-allocated counts");
        B200_RT_CHECK(b200sv_counts(dev_, w.data(), (int64_t)w.size(), (int64_t)GetDeviceShots(),
                                    &eigvals(0), stride1(eigvals), &counts(0), stride1(counts)));
    }
    void Counts(DataView<double, 1> &eigvals, DataView<int64_t, 1> &counts) override
    {
        counts_impl(eigvals, counts, {});
    }
    void PartialCounts(DataView<double, 1> &eigvals, DataView<int64_t, 1> &counts,
                       const std::vector<QubitIdType> &wires) override
    {
        const auto w = dev_wires(wires);
        RT_FAIL_IF(w.empty() || w.size() > GetNumQubits(), "Invalid number of wires");
        counts_impl(eigvals, counts, w);
    }
    auto Measure(QubitIdType wire, std::optional<int32_t> postselect = std::nullopt) -> Result override
    {
        const auto w = dev_wires({wire});
        int out = 0;
        B200_RT_CHECK(b200sv_measure(dev_, w[0], postselect ? (int)postselect.value() : -1, &out));
        return out ? &static_true_ : &static_false_;
    }

    auto PauliMeasure(const std::string &pauli_word, const std::vector<QubitIdType> &wires) -> Result override
    {
        const auto w = dev_wires(wires);
        int out = 0;
        B200_RT_CHECK(b200sv_pauli_measure(dev_, pauli_word.c_str(), w.data(), (int64_t)w.size(), -1, &out));
        return out ? &static_true_ : &static_false_;
    }

    // ---- differentiation (QuantumDevice.hpp:564-581) -----------------------------------------------
    void StartTapeRecording() override { B200_RT_CHECK(b200sv_start_tape_recording(dev_)); }
    void StopTapeRecording() override { B200_RT_CHECK(b200sv_stop_tape_recording(dev_)); }
    void Gradient(std::vector<DataView<double, 1>> &gradients,
                  const std::vector<size_t> &trainParams) override
    {
        std::vector<double *> rows;
        std::vector<int64_t> sizes, strides, train(trainParams.begin(), trainParams.end());
        for (auto &g : gradients) {
            rows.push_back(g.size() ? &g(0) : nullptr);
            sizes.push_back((int64_t)g.size());
            strides.push_back(stride1(g));
        }
        B200_RT_CHECK(b200sv_gradient(dev_, rows.data(), sizes.data(), strides.data(),
                                      (int64_t)rows.size(), train.data(), (int64_t)train.size()));
    }
};

} // namespace Catalyst::Runtime::Devices

GENERATE_DEVICE_FACTORY(B200StateVector, Catalyst::Runtime::Devices::B200StateVector);
