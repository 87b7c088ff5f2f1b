"""HARNESS (tests + bench e2e): drive a device plugin through the reference's own runtime C API.

`RuntimeDevice` loads oracle/_ref/librt_capi.so (the reference's runtime/lib/capi/RuntimeCAPI.cpp
compiled in place by oracle/build_ref.py) and issues exactly the calls a @qjit-compiled program
makes (reference: runtime/include/RuntimeCAPI.h:26-124; lowering conventions
mlir/lib/Quantum/Transforms/ConversionPatterns.cpp): __catalyst__rt__initialize ->
__catalyst__rt__device_init(lib, name, kwargs, shots, auto_qubit_management) ->
qubit_allocate_array -> __catalyst__qis__<Gate>(..., Modifiers*) -> NamedObs/Expval/Probs/... ->
qubit_release_array -> device_release -> finalize.  It exposes the same method set as
oracle.OracleDevice / capi.B200Device so the parity tests can run unchanged through the real
drop-in boundary (C++ vtable of librtd_b200sv.so).  Errors raised by the device would propagate
as C++ exceptions through ctypes, so only valid programs are driven this way.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RT_CAPI = os.path.join(ROOT, "oracle", "_ref", "librt_capi.so")
PLUGIN = os.path.join(ROOT, "catalyst_b200", "lib", "librtd_b200sv.so")
NULL_PLUGIN = os.path.join(ROOT, "oracle", "_ref", "librtd_null_qubit.so")

OBS_IDS = {"Identity": 0, "PauliX": 1, "PauliY": 2, "PauliZ": 3, "Hadamard": 4}


def available() -> bool:
    return os.path.exists(RT_CAPI) and os.path.exists(PLUGIN)


class Modifiers(C.Structure):  # runtime/include/Types.h:159-164
    _fields_ = [("adjoint", C.c_bool), ("num_controlled", C.c_size_t),
                ("controlled_wires", C.POINTER(C.c_void_p)), ("controlled_values", C.POINTER(C.c_bool))]


def _memref(ctype, rank):
    class M(C.Structure):  # runtime/include/Types.h:84-150
        _fields_ = [("allocated", C.POINTER(ctype)), ("aligned", C.POINTER(ctype)), ("offset", C.c_size_t),
                    ("sizes", C.c_size_t * rank), ("strides", C.c_size_t * rank)]
    return M


MR_d1, MR_d2 = _memref(C.c_double, 1), _memref(C.c_double, 2)
MR_i64 = _memref(C.c_int64, 1)
MR_i8 = _memref(C.c_int8, 1)


class PairCounts(C.Structure):
    _fields_ = [("eig", MR_d1), ("cnt", MR_i64)]


def _mr(arr, M, ctype):
    m = M()
    p = arr.ctypes.data_as(C.POINTER(ctype))
    m.allocated, m.aligned, m.offset = p, p, 0
    item = arr.itemsize if arr.dtype != np.complex128 else 16
    for i in range(arr.ndim):
        m.sizes[i] = arr.shape[i]
        m.strides[i] = arr.strides[i] // item
    return m


# gate symbol -> (n_params, n_wires); variadic ones handled separately
FIXED = {"PauliX": (0, 1), "PauliY": (0, 1), "PauliZ": (0, 1), "Hadamard": (0, 1), "S": (0, 1), "T": (0, 1),
         "PhaseShift": (1, 1), "RX": (1, 1), "RY": (1, 1), "RZ": (1, 1), "Rot": (3, 1), "CNOT": (0, 2),
         "CY": (0, 2), "CZ": (0, 2), "SWAP": (0, 2), "IsingXX": (1, 2), "IsingYY": (1, 2), "IsingXY": (1, 2),
         "IsingZZ": (1, 2), "SingleExcitation": (1, 2), "DoubleExcitation": (1, 4),
         "ControlledPhaseShift": (1, 2), "CRX": (1, 2), "CRY": (1, 2), "CRZ": (1, 2), "CRot": (3, 2),
         "CSWAP": (0, 3), "Toffoli": (0, 3), "ISWAP": (0, 2), "PSWAP": (1, 2)}


class _Syms:
    """Attribute access to the `__catalyst__*` symbols (avoids Python's private-name mangling)."""

    def __init__(self, cdll):
        object.__setattr__(self, "_cdll", cdll)

    def __getattr__(self, name):
        return getattr(self._cdll, "__catalyst__" + name)


class RuntimeDevice:
    _active = False

    def __init__(self, n, seed=None, shots=0, lib=PLUGIN, name="B200StateVector", **kwargs):
        assert not RuntimeDevice._active, "one runtime context at a time"
        self.L = _Syms(C.CDLL(RT_CAPI))
        L = self.L
        L.rt__qubit_allocate_array.restype = C.c_void_p
        L.rt__array_get_element_ptr_1d.restype = C.POINTER(C.c_void_p)
        L.rt__num_qubits.restype = C.c_int64
        L.qis__Measure.restype = C.POINTER(C.c_bool)
        L.qis__PauliMeasure.restype = C.POINTER(C.c_bool)
        L.qis__Expval.restype = C.c_double
        L.qis__Variance.restype = C.c_double
        for s in ("NamedObs", "HermitianObs", "TensorObs", "HamiltonianObs"):
            getattr(L, "qis__" + s).restype = C.c_int64
        self._seed = C.c_uint32(seed) if seed is not None else None
        L.rt__initialize(C.byref(self._seed) if seed is not None else None)
        RuntimeDevice._active = True
        self.lib, self.name = lib.encode(), name.encode()
        kwargs.setdefault("shots", shots)
        self.kwargs = str(dict(kwargs)).encode()
        self.shots = shots
        self.recorder = False
        self.n = 0
        self.qreg = None
        self._open(n)

    # -- lifecycle ----------------------------------------------------------------------------
    def _open(self, n):
        L = self.L
        L.rt__device_init(self.lib, self.name, self.kwargs, C.c_int64(self.shots), C.c_bool(False))
        self.qreg = C.c_void_p(L.rt__qubit_allocate_array(C.c_int64(n)))
        self.q = [C.c_void_p(L.rt__array_get_element_ptr_1d(self.qreg, C.c_int64(i))[0])
                  for i in range(n)]
        self.n = n

    def _close(self):
        if self.qreg is not None:
            self.L.rt__qubit_release_array(self.qreg)
            self.L.rt__device_release()
            self.qreg = None

    def reset(self, n=None):
        n = self.n if n is None else n
        self._close()
        self._open(n)

    def set_shots(self, shots):
        self.shots = shots
        n = self.n
        self._close()
        self._open(n)

    def close(self):
        if RuntimeDevice._active and getattr(self, "L", None) is not None:
            self._close()
            self.L.rt__finalize()
            RuntimeDevice._active = False
            self.L = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # pragma: no cover
            pass

    def num_qubits(self):
        return int(self.L.rt__num_qubits())

    # -- dynamic qubits (RuntimeCAPI.cpp:305-372): self.q[w] stays the id of device wire w ----------
    def allocate_qubit(self):
        self.L.rt__qubit_allocate.restype = C.c_void_p
        self.q.append(C.c_void_p(self.L.rt__qubit_allocate()))
        self.n += 1

    def release_qubit(self, wire):
        self.L.rt__qubit_release(self.q[wire])
        del self.q[wire]
        self.n -= 1

    def allocate_register(self, k):
        """qml.allocate(k) on a live device: a second QirArray; returns (handle, device wires)."""
        reg = C.c_void_p(self.L.rt__qubit_allocate_array(C.c_int64(k)))
        first = len(self.q)
        self.q += [C.c_void_p(self.L.rt__array_get_element_ptr_1d(reg, C.c_int64(i))[0]) for i in range(k)]
        self.n += k
        return reg, list(range(first, first + k))

    def release_register(self, reg, wires):
        self.L.rt__qubit_release_array(reg)
        for w in sorted(wires, reverse=True):
            del self.q[w]
        self.n -= len(wires)

    # -- gates --------------------------------------------------------------------------------
    def _mods(self, inverse, ctrl_wires, ctrl_values):
        if not inverse and not len(ctrl_wires):
            return None, None  # nullptr when there are no modifiers (ConversionPatterns.cpp:112-114)
        nc = len(ctrl_wires)
        cw = (C.c_void_p * max(nc, 1))(*[self.q[w] for w in ctrl_wires])
        cv = (C.c_bool * max(nc, 1))(*[bool(v) for v in ctrl_values])
        m = Modifiers(bool(inverse), nc, C.cast(cw, C.POINTER(C.c_void_p)), C.cast(cv, C.POINTER(C.c_bool)))
        return m, (cw, cv)

    def named_op(self, name, params=(), wires=(), inverse=False, ctrl_wires=(), ctrl_values=(),
                 pauli_word=None):
        m, keep = self._mods(inverse, list(ctrl_wires), list(ctrl_values))
        mp = C.byref(m) if m is not None else None
        qs = [self.q[w] for w in wires]
        ps = [C.c_double(p) for p in params]
        f = getattr(self.L, "qis__" + name)
        if name in FIXED:
            assert FIXED[name] == (len(ps), len(qs)), name
            f(*ps, *qs, mp)
        elif name == "Identity":
            f(mp, C.c_int64(len(qs)), *qs)
        elif name == "MultiRZ":
            f(ps[0], mp, C.c_int64(len(qs)), *qs)
        elif name == "GlobalPhase":
            f(ps[0], mp)
        elif name == "PCPhase":
            f(ps[0], C.c_int64(int(params[1])), mp, C.c_int64(len(qs)), *qs)
        elif name == "PauliRot":
            f(pauli_word.encode(), ps[0], mp, C.c_bool(True), C.c_int64(len(qs)), *qs)
        else:
            raise ValueError(name)

    def matrix_op(self, matrix, wires, inverse=False, ctrl_wires=(), ctrl_values=()):
        k = len(wires)
        a = np.ascontiguousarray(np.asarray(matrix, dtype=np.complex128).reshape(1 << k, 1 << k))
        mr = _mr(a.view(np.float64).reshape(1 << k, 2 << k), MR_d2, C.c_double)
        mr.sizes[1] = 1 << k
        mr.strides[0] = 1 << k
        m, keep = self._mods(inverse, list(ctrl_wires), list(ctrl_values))
        self.L.qis__QubitUnitary(C.byref(mr), C.byref(m) if m is not None else None,
                                             C.c_int64(k), *[self.q[w] for w in wires])

    def set_state(self, state, wires):
        a = np.ascontiguousarray(np.asarray(state, dtype=np.complex128))
        mr = _mr(a.view(np.float64), MR_d1, C.c_double)
        mr.sizes[0] = a.size
        self.L.qis__SetState(C.byref(mr), C.c_uint64(len(wires)), *[self.q[w] for w in wires])

    def set_basis_state(self, bits, wires):
        a = np.ascontiguousarray(np.asarray(list(bits), dtype=np.int8))
        mr = _mr(a, MR_i8, C.c_int8)
        self.L.qis__SetBasisState(C.byref(mr), C.c_uint64(len(wires)), *[self.q[w] for w in wires])

    # -- observables --------------------------------------------------------------------------
    def obs_named(self, name_or_id, wire):
        i = OBS_IDS[name_or_id] if isinstance(name_or_id, str) else int(name_or_id)
        w = wire if np.isscalar(wire) else wire[0]
        return int(self.L.qis__NamedObs(C.c_int64(i), self.q[w]))

    def obs_hermitian(self, matrix, wires):
        k = len(wires)
        a = np.ascontiguousarray(np.asarray(matrix, dtype=np.complex128).reshape(1 << k, 1 << k))
        mr = _mr(a.view(np.float64).reshape(1 << k, 2 << k), MR_d2, C.c_double)
        mr.sizes[1] = 1 << k
        mr.strides[0] = 1 << k
        return int(self.L.qis__HermitianObs(C.byref(mr), C.c_int64(k), *[self.q[w] for w in wires]))

    def obs_tensor(self, keys):
        return int(self.L.qis__TensorObs(C.c_int64(len(keys)), *[C.c_int64(k) for k in keys]))

    def obs_hamiltonian(self, coeffs, keys):
        c = np.ascontiguousarray(np.asarray(list(coeffs), dtype=np.float64))
        mr = _mr(c, MR_d1, C.c_double)
        return int(self.L.qis__HamiltonianObs(C.byref(mr), C.c_int64(len(keys)),
                                                          *[C.c_int64(k) for k in keys]))

    # -- measurements ---------------------------------------------------------------------------
    def expval(self, key):
        return float(self.L.qis__Expval(C.c_int64(key)))

    def var(self, key):
        return float(self.L.qis__Variance(C.c_int64(key)))

    def state(self, out=None):
        if out is None:
            out = np.empty(1 << self.n, dtype=np.complex128)
        mr = _mr(out.view(np.float64), MR_d1, C.c_double)
        mr.sizes[0] = out.size
        self.L.qis__State(C.byref(mr), C.c_int64(0))
        return out

    def probs(self, wires=None, out=None):
        wires = list(wires or [])
        k = len(wires) if wires else self.n
        if out is None:
            out = np.empty(1 << k, dtype=np.float64)
        mr = _mr(out, MR_d1, C.c_double)
        self.L.qis__Probs(C.byref(mr), C.c_int64(len(wires)), *[self.q[w] for w in wires])
        return out

    def sample(self, shots, wires=None, out=None):
        if shots != self.shots:
            raise ValueError("shots are fixed at device_init (RuntimeCAPI.cpp:1052-1056)")
        wires = list(wires or [])
        k = len(wires) if wires else self.n
        if out is None:
            out = np.empty((shots, k), dtype=np.float64)
        mr = _mr(out, MR_d2, C.c_double)
        self.L.qis__Sample(C.byref(mr), C.c_int64(len(wires)), *[self.q[w] for w in wires])
        return out

    def counts(self, shots, wires=None):
        if shots != self.shots:
            raise ValueError("shots are fixed at device_init")
        wires = list(wires or [])
        k = len(wires) if wires else self.n
        eig = np.empty(1 << k, dtype=np.float64)
        cnt = np.empty(1 << k, dtype=np.int64)
        pr = PairCounts(_mr(eig, MR_d1, C.c_double), _mr(cnt, MR_i64, C.c_int64))
        self.L.qis__Counts(C.byref(pr), C.c_int64(len(wires)), *[self.q[w] for w in wires])
        return eig, cnt

    def measure(self, wire, postselect=None):
        r = self.L.qis__Measure(self.q[wire], C.c_int32(-1 if postselect is None else postselect))
        return bool(r[0])

    def pauli_measure(self, word, wires, negated=False):
        # RuntimeCAPI.cpp:962-993: (pauliStr, negated, pauliStrAlt, negatedAlt, selectSwitch, numQubits, qubits...)
        r = self.L.qis__PauliMeasure(word.encode(), C.c_bool(negated), word.encode(), C.c_bool(negated),
                                     C.c_bool(True), C.c_int64(len(wires)), *[self.q[w] for w in wires])
        return bool(r[0])

    # -- gradient (lowering: mlir/lib/Gradient/Transforms/ConversionPatterns.cpp:198-274) ---------
    def start_tape(self):
        self.L.rt__toggle_recorder(C.c_bool(True))

    def stop_tape(self):
        self.L.rt__toggle_recorder(C.c_bool(False))

    def gradient(self, nobs, nparams, train=None):
        out = np.zeros((nobs, nparams if train is None else len(train)), dtype=np.float64)
        mrs = [_mr(out[k], MR_d1, C.c_double) for k in range(nobs)]
        ptrs = [C.byref(m) for m in mrs]
        if train is None:
            self.L.qis__Gradient(C.c_int64(nobs), *ptrs)
        else:
            t = np.ascontiguousarray(np.asarray(train, dtype=np.int64))
            tm = _mr(t, MR_i64, C.c_int64)
            self.L.qis__Gradient_params(C.byref(tm), C.c_int64(nobs), *ptrs)
        return out
