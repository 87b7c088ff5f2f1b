"""ctypes binding of the C ABI in include/b200sv.h (libb200sv.so).

``B200Device`` mirrors, method for method, the reference's plugin interface
(reference: runtime/include/QuantumDevice.hpp:51-582) with device wires 0..n-1, so tests read like
the reference's own device tests.  All arithmetic happens in the CUDA library; this module holds no
numerical code and there is no fallback: if the library or a CUDA device is missing, constructing a
device raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libb200sv.so")

OBS_IDS = {"Identity": 0, "PauliX": 1, "PauliY": 2, "PauliZ": 3, "Hadamard": 4, "Hermitian": 5}

# every symbol include/b200sv.h declares (checked by tests/test_abi.py)
SYMBOLS = [
    "b200sv_last_error", "b200sv_version", "b200sv_create", "b200sv_destroy",
    "b200sv_allocate_qubits", "b200sv_allocate_qubit", "b200sv_release_qubits", "b200sv_release_qubit",
    "b200sv_num_qubits",
    "b200sv_set_shots", "b200sv_get_shots", "b200sv_set_prng", "b200sv_seed", "b200sv_prng_peek",
    "b200sv_named_operation", "b200sv_matrix_operation", "b200sv_set_state",
    "b200sv_set_basis_state", "b200sv_observable", "b200sv_tensor_observable",
    "b200sv_hamiltonian_observable", "b200sv_expval", "b200sv_var", "b200sv_state", "b200sv_probs",
    "b200sv_sample", "b200sv_counts", "b200sv_measure", "b200sv_pauli_measure", "b200sv_start_tape_recording",
    "b200sv_stop_tape_recording", "b200sv_tape_num_params", "b200sv_tape_num_observables",
    "b200sv_gradient", "b200sv_get_stats", "b200sv_reset_stats", "b200sv_set_timing",
    "b200sv_flush", "b200sv_fp64_peak", "b200sv_stats_json", "b200sv_synchronize", "b200sv_jit_wait", "b200sv_device_pointer",
    "b200sv_nccl_unique_id", "b200sv_world_size", "b200sv_rank", "b200sv_layout",
]


class Stats(C.Structure):
    _fields_ = [("gates_enqueued", C.c_int64), ("kernel_ops", C.c_int64),
                ("tile_passes", C.c_int64), ("other_launches", C.c_int64),
                ("gate_device_ms", C.c_double), ("algorithmic_bytes", C.c_double),
                ("rounds", C.c_int64), ("exchanges", C.c_int64), ("exchange_bytes", C.c_double),
                ("jit_passes", C.c_int64), ("jit_compiled", C.c_int64), ("jit_pending", C.c_int64),
                ("jit_compile_ms", C.c_double), ("p2p_exchanges", C.c_int64),
                ("relabelled_swaps", C.c_int64), ("swap_device_ms", C.c_double), ("fp64_ops", C.c_double),
                ("jit_disk_hits", C.c_int64), ("swap_push_exchanges", C.c_int64), ("d2h_bytes", C.c_double),
                ("h2d_bytes", C.c_double)]


_lib = None
_NAME_BYTES = {}


def lib():
    """Load libb200sv.so (built in-tree by catalyst_b200/build.py).  Raises if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -m catalyst_b200.build` "
                               "(there is no fallback implementation)")
        L = C.CDLL(LIB_PATH)
        L.b200sv_last_error.restype = C.c_char_p
        L.b200sv_version.restype = C.c_char_p
        L.b200sv_create.restype = C.c_void_p
        L.b200sv_create.argtypes = [C.c_char_p]
        L.b200sv_destroy.argtypes = [C.c_void_p]
        L.b200sv_device_pointer.restype = C.c_void_p
        L.b200sv_device_pointer.argtypes = [C.c_void_p]
        L.b200sv_prng_peek.restype = C.c_uint32
        L.b200sv_layout.restype = C.c_int64
        L.b200sv_layout.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.c_int64]
        L.b200sv_named_operation.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_double), C.c_int64,
                                             C.POINTER(C.c_int64), C.c_int64, C.c_int, C.POINTER(C.c_int64),
                                             C.POINTER(C.c_uint8), C.c_int64, C.c_char_p]
        L.b200sv_world_size.argtypes = [C.c_void_p]
        L.b200sv_rank.argtypes = [C.c_void_p]
        for name in ("b200sv_num_qubits", "b200sv_get_shots", "b200sv_observable",
                     "b200sv_tensor_observable", "b200sv_hamiltonian_observable",
                     "b200sv_tape_num_params", "b200sv_tape_num_observables"):
            getattr(L, name).restype = C.c_int64
        _lib = L
    return _lib


def nccl_unique_id() -> str:
    """256 hex digits of a fresh ncclUniqueId (rank 0 creates it, the host program distributes it)."""
    L = lib()
    buf = C.create_string_buffer(257)
    if L.b200sv_nccl_unique_id(buf) != 0:
        raise B200Error(L.b200sv_last_error().decode())
    return buf.value.decode()


def run_spmd(world: int, fn, group: str | None = None, **kwargs):
    """Run ``fn(device, rank)`` on `world` ranks as threads of this process (comm 'local:<group>'):
    the single-process form of the sharded device, also usable with all ranks on ONE GPU.
    Returns the list of results in rank order; re-raises the first failure."""
    import threading
    import uuid
    group = group or ("g" + uuid.uuid4().hex[:12])
    out, err = [None] * world, [None] * world

    def work(r):
        d = None
        try:
            kw = dict(kwargs)
            devices = kw.pop("devices", None)   # one CUDA ordinal per rank: ranks on different GPUs
            if devices is not None:
                kw["device"] = int(devices[r])
            if world > 1:
                kw.update(world_size=world, rank=r, comm=f"local:{group}")
            d = B200Device(**kw)
            out[r] = fn(d, r)
        except BaseException as e:  # noqa: BLE001
            err[r] = e
        finally:
            if d is not None:
                d.close()

    th = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for e in err:
        if e is not None:
            raise e
    return out


def _i64(a):
    return np.ascontiguousarray(np.asarray(list(a), dtype=np.int64))


def _p(a, ty):
    return a.ctypes.data_as(C.POINTER(ty))


class B200Error(RuntimeError):
    pass


class B200Device:
    """The CUDA device with the QuantumDevice method set (device wires)."""

    def __init__(self, n: int | None = None, seed: int | None = None, **kwargs):
        self.L = lib()
        self._named = self.L.b200sv_named_operation
        self.h = C.c_void_p(self.L.b200sv_create(str(dict(kwargs)).encode()))
        if not self.h:
            raise B200Error(self.L.b200sv_last_error().decode())
        self.n = 0
        if seed is not None:
            self.seed(seed)
        if n is not None:
            self.allocate_qubits(n)

    def close(self):
        if getattr(self, "h", None):
            self.L.b200sv_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # pragma: no cover
            pass

    def _chk(self, rc):
        if rc != 0:
            raise B200Error(self.L.b200sv_last_error().decode())

    # -- lifecycle (QuantumDevice.hpp:88-192) ---------------------------------------------------
    def allocate_qubits(self, n):
        self._chk(self.L.b200sv_allocate_qubits(self.h, C.c_int64(n)))
        self.n = n

    def allocate_qubit(self):
        self._chk(self.L.b200sv_allocate_qubit(self.h))
        self.n = int(self.L.b200sv_num_qubits(self.h))

    def release_qubits(self):
        self._chk(self.L.b200sv_release_qubits(self.h))
        self.n = 0

    def release_qubit(self, wire):
        self._chk(self.L.b200sv_release_qubit(self.h, C.c_int64(wire)))
        self.n = int(self.L.b200sv_num_qubits(self.h))

    def reset(self, n=None):
        n = self.n if n is None else n
        self.release_qubits()
        self.allocate_qubits(n)

    def num_qubits(self):
        return int(self.L.b200sv_num_qubits(self.h))

    def set_shots(self, shots):
        self._chk(self.L.b200sv_set_shots(self.h, C.c_int64(shots)))

    def get_shots(self):
        return int(self.L.b200sv_get_shots(self.h))

    def seed(self, s):
        self._chk(self.L.b200sv_seed(self.h, C.c_uint32(s)))

    def gen_peek(self) -> int:
        return int(self.L.b200sv_prng_peek(self.h))

    # -- gates (QuantumDevice.hpp:215-284) ------------------------------------------------------
    def named_op(self, name, params=(), wires=(), inverse=False, ctrl_wires=(), ctrl_values=(),
                 pauli_word=None):
        # the hot host call (one per gate): plain ctypes arrays, no numpy round trip
        bname = _NAME_BYTES.get(name)
        if bname is None:
            bname = _NAME_BYTES[name] = name.encode()
        np_, nw, nc = len(params), len(wires), len(ctrl_wires)
        p = (C.c_double * np_)(*params) if np_ else None
        w = (C.c_int64 * nw)(*wires) if nw else None
        if nc:
            cw = (C.c_int64 * nc)(*ctrl_wires)
            cv = (C.c_uint8 * nc)(*[1 if v else 0 for v in ctrl_values])
        else:
            cw = cv = None
        rc = self._named(self.h, bname, p, np_, w, nw, 1 if inverse else 0, cw, cv, nc,
                         pauli_word.encode() if pauli_word else None)
        if rc != 0:
            raise B200Error(self.L.b200sv_last_error().decode())

    def matrix_op(self, matrix, wires, inverse=False, ctrl_wires=(), ctrl_values=()):
        m = np.ascontiguousarray(np.asarray(matrix, dtype=np.complex128).reshape(-1))
        w, cw = _i64(wires), _i64(ctrl_wires)
        cv = np.ascontiguousarray(np.asarray(list(ctrl_values), dtype=np.uint8))
        if m.size != (1 << len(w)) ** 2:
            raise B200Error("Invalid given matrix size")
        self._chk(self.L.b200sv_matrix_operation(
            self.h, _p(m.view(np.float64), C.c_double), _p(w, C.c_int64), C.c_int64(len(w)),
            int(bool(inverse)), _p(cw, C.c_int64), _p(cv, C.c_uint8), C.c_int64(len(cw))))

    def set_state(self, state, wires):
        s = np.ascontiguousarray(np.asarray(state, dtype=np.complex128))
        w = _i64(wires)
        if s.size != 1 << len(w):
            raise B200Error("Invalid state size: must be 2^len(wires)")
        self._chk(self.L.b200sv_set_state(self.h, _p(s.view(np.float64), C.c_double), C.c_int64(1),
                                          _p(w, C.c_int64), C.c_int64(len(w))))

    def set_basis_state(self, bits, wires):
        b = np.ascontiguousarray(np.asarray(list(bits), dtype=np.int8))
        w = _i64(wires)
        if len(b) != len(w):
            raise B200Error("BasisState parameter and wires must be of equal length.")
        self._chk(self.L.b200sv_set_basis_state(self.h, _p(b, C.c_int8), C.c_int64(1),
                                                _p(w, C.c_int64), C.c_int64(len(w))))

    # -- observables (QuantumDevice.hpp:317-353) ------------------------------------------------
    def _key(self, k):
        if k < 0:
            raise B200Error(self.L.b200sv_last_error().decode())
        return int(k)

    def obs_named(self, name_or_id, wire):
        i = OBS_IDS[name_or_id] if isinstance(name_or_id, str) else int(name_or_id)
        w = _i64([wire] if np.isscalar(wire) else wire)
        return self._key(self.L.b200sv_observable(self.h, i, None, C.c_int64(0), _p(w, C.c_int64),
                                                  C.c_int64(len(w))))

    def obs_hermitian(self, matrix, wires):
        m = np.ascontiguousarray(np.asarray(matrix, dtype=np.complex128).reshape(-1))
        w = _i64(wires)
        return self._key(self.L.b200sv_observable(self.h, 5, _p(m.view(np.float64), C.c_double),
                                                  C.c_int64(m.size), _p(w, C.c_int64),
                                                  C.c_int64(len(w))))

    def obs_tensor(self, keys):
        k = _i64(keys)
        return self._key(self.L.b200sv_tensor_observable(self.h, _p(k, C.c_int64), C.c_int64(len(k))))

    def obs_hamiltonian(self, coeffs, keys):
        c = np.ascontiguousarray(np.asarray(list(coeffs), dtype=np.float64))
        k = _i64(keys)
        if len(c) != len(k):
            raise B200Error("Invalid size for the list of coefficients and observables")
        return self._key(self.L.b200sv_hamiltonian_observable(self.h, _p(c, C.c_double),
                                                              _p(k, C.c_int64), C.c_int64(len(k))))

    # -- measurements (QuantumDevice.hpp:234, 384-518) ------------------------------------------
    def expval(self, key):
        out = C.c_double()
        self._chk(self.L.b200sv_expval(self.h, C.c_int64(key), C.byref(out)))
        return out.value

    def var(self, key):
        out = C.c_double()
        self._chk(self.L.b200sv_var(self.h, C.c_int64(key), C.byref(out)))
        return out.value

    def state(self, out=None):
        if out is None:
            out = np.empty(1 << self.n, dtype=np.complex128)
        self._chk(self.L.b200sv_state(self.h, _p(out.view(np.float64), C.c_double), C.c_int64(1)))
        return out

    def probs(self, wires=None, out=None):
        w = _i64(wires or [])
        k = len(w) if len(w) else self.n
        if out is None:
            out = np.empty(1 << k, dtype=np.float64)
        self._chk(self.L.b200sv_probs(self.h, _p(w, C.c_int64), C.c_int64(len(w)),
                                      _p(out, C.c_double), C.c_int64(1)))
        return out

    def sample(self, shots, wires=None):
        w = _i64(wires or [])
        k = len(w) if len(w) else self.n
        out = np.empty((shots, k), dtype=np.float64)
        self._chk(self.L.b200sv_sample(self.h, _p(w, C.c_int64), C.c_int64(len(w)),
                                       C.c_int64(shots), _p(out, C.c_double), C.c_int64(k),
                                       C.c_int64(1)))
        return out

    def counts(self, shots, wires=None):
        w = _i64(wires or [])
        k = len(w) if len(w) else self.n
        eig = np.empty(1 << k, dtype=np.float64)
        cnt = np.empty(1 << k, dtype=np.int64)
        self._chk(self.L.b200sv_counts(self.h, _p(w, C.c_int64), C.c_int64(len(w)),
                                       C.c_int64(shots), _p(eig, C.c_double), C.c_int64(1),
                                       _p(cnt, C.c_int64), C.c_int64(1)))
        return eig, cnt

    def measure(self, wire, postselect=None):
        out = C.c_int()
        self._chk(self.L.b200sv_measure(self.h, C.c_int64(wire),
                                        -1 if postselect is None else int(postselect),
                                        C.byref(out)))
        return bool(out.value)

    def pauli_measure(self, word, wires, postselect=None):
        """PauliMeasure (QuantumDevice.hpp:531): False <-> eigenvalue +1 of the Pauli product."""
        out = C.c_int()
        w = (C.c_int64 * len(wires))(*wires)
        self._chk(self.L.b200sv_pauli_measure(self.h, word.encode(), w, C.c_int64(len(wires)),
                                              -1 if postselect is None else int(postselect), C.byref(out)))
        return bool(out.value)

    # -- tape / gradient (QuantumDevice.hpp:564-581) --------------------------------------------
    def start_tape(self):
        self._chk(self.L.b200sv_start_tape_recording(self.h))

    def stop_tape(self):
        self._chk(self.L.b200sv_stop_tape_recording(self.h))

    def gradient(self, train=None):
        t = _i64(train or [])
        P = int(self.L.b200sv_tape_num_params(self.h))
        K = int(self.L.b200sv_tape_num_observables(self.h))
        T = len(t) if len(t) else P
        out = np.zeros((K, T), dtype=np.float64)
        if K * T:
            rows = (C.POINTER(C.c_double) * K)(*[_p(out[k], C.c_double) for k in range(K)])
            sizes = _i64([T] * K)
            strides = _i64([1] * K)
            self._chk(self.L.b200sv_gradient(self.h, rows, _p(sizes, C.c_int64),
                                             _p(strides, C.c_int64), C.c_int64(K),
                                             _p(t, C.c_int64), C.c_int64(len(t))))
        return out

    # -- instrumentation ------------------------------------------------------------------------
    def stats(self) -> Stats:
        s = Stats()
        self._chk(self.L.b200sv_get_stats(self.h, C.byref(s)))
        return s

    def reset_stats(self):
        self._chk(self.L.b200sv_reset_stats(self.h))

    def set_timing(self, on=True):
        self._chk(self.L.b200sv_set_timing(self.h, int(bool(on))))

    def flush(self):
        self._chk(self.L.b200sv_flush(self.h))

    def fp64_peak(self) -> float:
        """Measured thread-level FP64 FMA instructions per second of this GPU (microbenchmark)."""
        out = C.c_double()
        self._chk(self.L.b200sv_fp64_peak(self.h, C.byref(out)))
        return out.value

    def stats_json(self) -> dict:
        import json
        buf = C.create_string_buffer(4096)
        self.L.b200sv_stats_json.restype = C.c_int64
        if self.L.b200sv_stats_json(self.h, buf, C.c_int64(4096)) < 0:
            raise B200Error(self.L.b200sv_last_error().decode())
        return json.loads(buf.value.decode())

    def synchronize(self):
        self._chk(self.L.b200sv_synchronize(self.h))

    def jit_wait(self):
        """Block until no specialised tile pass is being compiled in the background."""
        self._chk(self.L.b200sv_jit_wait(self.h))

    def world_size(self) -> int:
        return int(self.L.b200sv_world_size(self.h))

    def rank(self) -> int:
        return int(self.L.b200sv_rank(self.h))

    def layout(self):
        buf = (C.c_int32 * 64)()
        k = int(self.L.b200sv_layout(self.h, buf, C.c_int64(64)))
        return [int(buf[i]) for i in range(k)]

    def device_pointer(self) -> int:
        return int(self.L.b200sv_device_pointer(self.h) or 0)
