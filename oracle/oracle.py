"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes driver for the CPU restatement in ``sv_oracle.hpp`` (see that file's header for what it
restates and how it is pinned).  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import this module; the product
(``catalyst_b200/``) never does.

The class mirrors the method set of the reference plugin interface
(reference: runtime/include/QuantumDevice.hpp:51-582) with device wires 0..n-1, so a test can run
the same gate sequence on the oracle and on the CUDA device.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libsv_oracle.so")

OBS_IDS = {"Identity": 0, "PauliX": 1, "PauliY": 2, "PauliZ": 3, "Hadamard": 4}


def build(force: bool = False) -> str:
    """Compile the oracle (g++ -O3 -fopenmp).  Building the checker is not using it."""
    src = [os.path.join(_HERE, "sv_oracle_capi.cpp"), os.path.join(_HERE, "sv_oracle.hpp")]
    if (not force and os.path.exists(_LIB_PATH)
            and all(os.path.getmtime(_LIB_PATH) >= os.path.getmtime(s) for s in src)):
        return _LIB_PATH
    os.makedirs(os.path.dirname(_LIB_PATH), exist_ok=True)
    cmd = ["g++", "-std=c++17", "-O3", "-march=x86-64-v3", "-fopenmp", "-shared", "-fPIC",
           "-o", _LIB_PATH, src[0]]
    subprocess.check_call(cmd)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = _LIB_PATH if os.path.exists(_LIB_PATH) else build()
        L = C.CDLL(path)
        L.orc_last_error.restype = C.c_char_p
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_int]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_gen_peek.restype = C.c_uint32
        for name in ("orc_obs_named", "orc_obs_hermitian", "orc_obs_tensor", "orc_obs_hamiltonian"):
            getattr(L, name).restype = C.c_int64
        _lib = L
    return _lib


def _i64(a):
    return np.ascontiguousarray(np.asarray(list(a), dtype=np.int64))


def _p(a, ty):
    return a.ctypes.data_as(C.POINTER(ty))


class OracleDevice:
    """CPU oracle with the QuantumDevice method set (device wires)."""

    def __init__(self, n: int, seed: int | None = None, seed_every_measurement: bool = False):
        self.L = lib()
        self.h = C.c_void_p(self.L.orc_create(int(n)))
        if not self.h:
            raise RuntimeError(self.L.orc_last_error().decode())
        self.n = n
        if seed is not None:
            self.L.orc_seed(self.h, C.c_uint32(seed))
        self.L.orc_set_seed_every_measurement(self.h, int(seed_every_measurement))

    def __del__(self):
        try:
            self.L.orc_destroy(self.h)
        except Exception:  # pragma: no cover
            pass

    def _chk(self, rc):
        if rc != 0:
            raise RuntimeError(self.L.orc_last_error().decode())

    # -- lifecycle ----------------------------------------------------------------------------
    def reset(self, n=None):
        self.n = self.n if n is None else n
        self._chk(self.L.orc_reset(self.h, int(self.n)))

    def allocate_qubit(self):
        """AllocateQubit (QuantumDevice.hpp:122): appends one wire in |0>."""
        self._chk(self.L.orc_allocate_qubit(self.h))
        self.n += 1

    def release_qubit(self, wire):
        """ReleaseQubit (QuantumDevice.hpp:144): the wire leaves the register, later wires move down."""
        self._chk(self.L.orc_release_qubit(self.h, C.c_int64(wire)))
        self.n -= 1

    def seed(self, s):
        self.L.orc_seed(self.h, C.c_uint32(s))

    def set_shots(self, shots):
        """SetDeviceShots (QuantumDevice.hpp:164): > 0 turns Expval / Var / Probs into sampled estimates."""
        self.L.orc_set_shots(self.h, C.c_int64(shots))

    def gen_peek(self) -> int:
        return int(self.L.orc_gen_peek(self.h))

    def set_mcmc(self, on=True, num_burnin=100, kernel_name="Local"):
        """Device kwargs mcmc / num_burnin / kernel_name of lightning.qubit: Sample / Counts / finite-shot
        estimates draw from a Metropolis chain instead of the alias sampler (parity unpinned, see sv_oracle.hpp)."""
        self._chk(self.L.orc_set_mcmc(self.h, int(bool(on)), C.c_int64(num_burnin), kernel_name.encode()))

    # -- gates --------------------------------------------------------------------------------
    def named_op(self, name, params=(), wires=(), inverse=False, ctrl_wires=(), ctrl_values=(),
                 pauli_word=None):
        p = np.ascontiguousarray(np.asarray(list(params), dtype=np.float64))
        w, cw = _i64(wires), _i64(ctrl_wires)
        cv = np.ascontiguousarray(np.asarray(list(ctrl_values), dtype=np.uint8))
        self._chk(self.L.orc_named_op(
            self.h, name.encode(), _p(p, C.c_double), len(p), _p(w, C.c_int64), len(w),
            int(bool(inverse)), _p(cw, C.c_int64), _p(cv, C.c_uint8), len(cw),
            pauli_word.encode() if pauli_word else None))

    def matrix_op(self, matrix, wires, inverse=False, ctrl_wires=(), ctrl_values=()):
        m = np.ascontiguousarray(np.asarray(matrix, dtype=np.complex128).reshape(-1))
        w, cw = _i64(wires), _i64(ctrl_wires)
        cv = np.ascontiguousarray(np.asarray(list(ctrl_values), dtype=np.uint8))
        self._chk(self.L.orc_matrix_op(self.h, _p(m.view(np.float64), C.c_double),
                                       _p(w, C.c_int64), len(w), int(bool(inverse)),
                                       _p(cw, C.c_int64), _p(cv, C.c_uint8), len(cw)))

    def set_state(self, state, wires):
        s = np.ascontiguousarray(np.asarray(state, dtype=np.complex128))
        w = _i64(wires)
        self._chk(self.L.orc_set_state(self.h, _p(s.view(np.float64), C.c_double),
                                       _p(w, C.c_int64), len(w)))

    def set_basis_state(self, bits, wires):
        b = np.ascontiguousarray(np.asarray(list(bits), dtype=np.int8))
        w = _i64(wires)
        self._chk(self.L.orc_set_basis_state(self.h, _p(b, C.c_int8), _p(w, C.c_int64), len(w)))

    # -- observables ----------------------------------------------------------------------------
    def _key(self, k):
        if k < 0:
            raise RuntimeError(self.L.orc_last_error().decode())
        return int(k)

    def obs_named(self, name_or_id, wire):
        i = OBS_IDS[name_or_id] if isinstance(name_or_id, str) else int(name_or_id)
        w = _i64([wire] if np.isscalar(wire) else wire)
        return self._key(self.L.orc_obs_named(self.h, i, _p(w, C.c_int64), len(w)))

    def obs_hermitian(self, matrix, wires):
        m = np.ascontiguousarray(np.asarray(matrix, dtype=np.complex128).reshape(-1))
        w = _i64(wires)
        return self._key(self.L.orc_obs_hermitian(self.h, _p(m.view(np.float64), C.c_double),
                                                  _p(w, C.c_int64), len(w)))

    def obs_tensor(self, keys):
        k = _i64(keys)
        return self._key(self.L.orc_obs_tensor(self.h, _p(k, C.c_int64), len(k)))

    def obs_hamiltonian(self, coeffs, keys):
        c = np.ascontiguousarray(np.asarray(list(coeffs), dtype=np.float64))
        k = _i64(keys)
        return self._key(self.L.orc_obs_hamiltonian(self.h, _p(c, C.c_double), _p(k, C.c_int64),
                                                    len(k)))

    # -- measurements ---------------------------------------------------------------------------
    def expval(self, key):
        out = C.c_double()
        self._chk(self.L.orc_expval(self.h, C.c_int64(key), C.byref(out)))
        return out.value

    def var(self, key):
        out = C.c_double()
        self._chk(self.L.orc_var(self.h, C.c_int64(key), C.byref(out)))
        return out.value

    def state(self):
        out = np.empty(1 << self.n, dtype=np.complex128)
        self._chk(self.L.orc_state(self.h, _p(out.view(np.float64), C.c_double)))
        return out

    def probs(self, wires=None):
        w = _i64(wires or [])
        k = len(w) if len(w) else self.n
        out = np.empty(1 << k, dtype=np.float64)
        self._chk(self.L.orc_probs(self.h, _p(w, C.c_int64), len(w), _p(out, C.c_double)))
        return out

    def sample(self, shots, wires=None):
        w = _i64(wires or [])
        k = len(w) if len(w) else self.n
        out = np.empty((shots, k), dtype=np.float64)
        self._chk(self.L.orc_sample(self.h, _p(w, C.c_int64), len(w), C.c_int64(shots),
                                    _p(out, C.c_double)))
        return out

    def counts(self, shots, wires=None):
        w = _i64(wires or [])
        k = len(w) if len(w) else self.n
        eig = np.empty(1 << k, dtype=np.float64)
        cnt = np.empty(1 << k, dtype=np.int64)
        self._chk(self.L.orc_counts(self.h, _p(w, C.c_int64), len(w), C.c_int64(shots),
                                    _p(eig, C.c_double), _p(cnt, C.c_int64)))
        return eig, cnt

    def measure(self, wire, postselect=None):
        out = C.c_int()
        self._chk(self.L.orc_measure(self.h, C.c_int64(wire),
                                     -1 if postselect is None else int(postselect), C.byref(out)))
        return bool(out.value)

    def pauli_measure(self, word, wires, postselect=None):
        out = C.c_int()
        w = (C.c_int64 * len(wires))(*wires)
        self._chk(self.L.orc_pauli_measure(self.h, word.encode(), w, len(wires),
                                           -1 if postselect is None else int(postselect), C.byref(out)))
        return bool(out.value)

    # -- tape / gradient ------------------------------------------------------------------------
    def start_tape(self):
        self._chk(self.L.orc_start_tape(self.h))

    def stop_tape(self):
        self._chk(self.L.orc_stop_tape(self.h))

    def gradient(self, train=None):
        t = _i64(train or [])
        P = self.L.orc_tape_num_params(self.h)
        K = self.L.orc_tape_num_obs(self.h)
        T = len(t) if len(t) else P
        out = np.zeros((K, T), dtype=np.float64)
        if K * T:
            self._chk(self.L.orc_gradient(self.h, _p(t, C.c_int64), len(t), _p(out, C.c_double)))
        return out


def num_threads() -> int:
    return int(lib().orc_num_threads())


def set_num_threads(t: int) -> None:
    lib().orc_set_num_threads(int(t))
