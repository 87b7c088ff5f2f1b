"""Python half of the drop-in boundary: the `qml.devices.Device` a user passes to `qml.qnode` so
that `@qjit` bakes (library path, factory name, kwargs) into the `quantum.device` op.

Contract (reference: doc/dev/custom_devices.rst:119-168; reader
frontend/catalyst/device/qjit_device.py:152-205): a `config_filepath` pointing at a schema-3 TOML,
a static `get_c_interface()` returning (factory identifier, absolute .so path), and an optional
`device_kwargs` dict that reaches the factory as `str(dict)`.  PennyLane is not installed in the
build container, so the class is only defined where `pennylane` imports; `c_interface()` below is
usable without it.
"""
from __future__ import annotations

import os

_HERE = os.path.dirname(os.path.abspath(__file__))
CONFIG_FILEPATH = os.path.join(_HERE, "b200sv.toml")
PLUGIN_PATH = os.path.join(_HERE, "lib", "librtd_b200sv.so")
FACTORY_IDENTIFIER = "B200StateVector"  # runtime dlsym()s FACTORY_IDENTIFIER + "Factory"


# ---- interpreted execution: walk a tape into the device ------------------------------
# `qml.devices.Device.execute(circuits, execution_config)` receives QuantumScript objects.  The walker below is
# duck-typed against that protocol (attributes only, no PennyLane import), so it can be exercised with a stand-in
# tape class where PennyLane is absent (tests/test_device_execute.py): operations carry `name`, `parameters`,
# `wires` and, for `Adjoint(..)` / `C(..)` wrappers, `base` / `control_wires` / `control_values`; measurements are
# recognised by class name (`ExpectationMP`, `VarianceMP`, `ProbabilityMP`, `StateMP`, `SampleMP`, `CountsMP`)
# with `obs` / `wires`; observables are named Paulis, `Hermitian`, `Prod` / `Tensor` (operands), `SProd`
# (scalar, base) and `Sum` / `Hamiltonian` / `LinearCombination` (`terms()` -> coeffs, ops).
_NAMED_OBS = ("Identity", "PauliX", "PauliY", "PauliZ", "Hadamard")
_PRODUCTS = ("Prod", "Tensor")
_SUMS = ("Sum", "Hamiltonian", "LinearCombination")


def _unwrap(op):
    """(base op, inverse, ctrl_wires, ctrl_values) of an operation that may be wrapped in Adjoint / Controlled."""
    inverse, cw, cv = False, [], []
    while True:
        name = op.name
        if name.startswith("Adjoint(") and hasattr(op, "base"):
            inverse = not inverse
            op = op.base
        elif (name.startswith("C(") or name == "Controlled") and hasattr(op, "base"):
            cw += list(op.control_wires)
            vals = getattr(op, "control_values", None)
            cv += [bool(v) for v in (vals if vals is not None else [True] * len(op.control_wires))]
            op = op.base
        else:
            return op, inverse, cw, cv


def _build_observable(dev, obs, wire_of):
    name = obs.name
    if name in _NAMED_OBS:
        return dev.obs_named(name, [wire_of(w) for w in obs.wires] if name == "Identity" else wire_of(list(obs.wires)[0]))
    if name == "Hermitian":
        return dev.obs_hermitian(obs.parameters[0], [wire_of(w) for w in obs.wires])
    if name in _PRODUCTS:
        return dev.obs_tensor([_build_observable(dev, o, wire_of) for o in obs.operands])
    if name == "SProd":
        return dev.obs_hamiltonian([float(obs.scalar)], [_build_observable(dev, obs.base, wire_of)])
    if name in _SUMS:
        coeffs, ops = obs.terms()
        return dev.obs_hamiltonian([float(c) for c in coeffs], [_build_observable(dev, o, wire_of) for o in ops])
    raise NotImplementedError(f"observable {name} is not supported by b200.statevector")


def execute_tape(tape, make_device, wires=None):
    """Run one tape on a fresh device object from `make_device(n, shots)` (capi.B200Device in production) and
    return its measurement results in order.  `wires`: the device's wire labels (default: sorted tape wires)."""
    import numpy as np
    labels = list(wires) if wires is not None else sorted(set(tape.wires))
    index = {w: i for i, w in enumerate(labels)}

    def wire_of(w):
        return index[w]

    shots = getattr(getattr(tape, "shots", None), "total_shots", None) or 0
    dev = make_device(len(labels), int(shots))
    try:
        for op in tape.operations:
            base, inverse, cw, cv = _unwrap(op)
            ws = [wire_of(w) for w in base.wires]
            cws = [wire_of(w) for w in cw]
            if base.name in ("QubitUnitary", "ControlledQubitUnitary"):
                dev.matrix_op(np.asarray(base.parameters[0], dtype=np.complex128), ws, inverse=inverse, ctrl_wires=cws,
                              ctrl_values=cv)
            elif base.name == "StatePrep":
                dev.set_state(np.asarray(base.parameters[0], dtype=np.complex128), ws)
            elif base.name == "BasisState":
                dev.set_basis_state([int(b) for b in base.parameters[0]], ws)
            else:
                params = [float(p) for p in base.parameters]
                word = getattr(base, "hyperparameters", {}).get("pauli_word") if base.name == "PauliRot" else None
                dev.named_op(base.name, params, ws, inverse=inverse, ctrl_wires=cws, ctrl_values=cv, pauli_word=word)
        out = []
        for m in tape.measurements:
            kind = type(m).__name__
            mw = [wire_of(w) for w in (m.wires if m.wires is not None else [])]
            if kind == "ExpectationMP":
                out.append(dev.expval(_build_observable(dev, m.obs, wire_of)))
            elif kind == "VarianceMP":
                out.append(dev.var(_build_observable(dev, m.obs, wire_of)))
            elif kind == "ProbabilityMP":
                out.append(dev.probs(mw or None))
            elif kind == "StateMP":
                out.append(dev.state())
            elif kind == "SampleMP":
                out.append(dev.sample(int(shots), mw or None))
            elif kind == "CountsMP":
                out.append(dev.counts(int(shots), mw or None))
            else:
                raise NotImplementedError(f"measurement process {kind} is not supported by b200.statevector")
        return tuple(out) if len(out) != 1 else out[0]
    finally:
        close = getattr(dev, "close", None)
        if close:
            close()


def c_interface():
    """(name, path) exactly as `get_c_interface()` must return them."""
    if not os.path.exists(PLUGIN_PATH):
        raise RuntimeError(f"{PLUGIN_PATH} is missing: run `python -m catalyst_b200.build`")
    return FACTORY_IDENTIFIER, PLUGIN_PATH


try:  # pragma: no cover - exercised only where PennyLane is installed
    import pennylane as qml

    class B200StateVectorDevice(qml.devices.Device):
        """`qml.device`-style handle: `dev = B200StateVectorDevice(wires=30, shots=None)`."""

        name = "b200.statevector"
        config_filepath = CONFIG_FILEPATH

        def __init__(self, wires=None, shots=None, tile_bits=None, fuse=True, cuda_device=None,
                     world_size=1, rank=0, comm=None, **options):
            super().__init__(wires=wires, shots=shots)
            # further factory kwargs travel unchanged (include/b200sv.h: mcmc / num_burnin / kernel_name, gather,
            # dist_sampler, release_entangled, jit, ...)
            self.device_kwargs = {"fuse": bool(fuse), **options}
            if tile_bits is not None:
                self.device_kwargs["tile_bits"] = int(tile_bits)
            if cuda_device is not None:
                self.device_kwargs["device"] = int(cuda_device)
            if world_size > 1:  # SPMD sharding over GPUs: see include/b200sv.h, "sharding"
                self.device_kwargs.update(world_size=int(world_size), rank=int(rank), comm=str(comm))

        @staticmethod
        def get_c_interface():
            return c_interface()

        def execute(self, circuits, execution_config=None):
            """Interpreted execution (outside @qjit): every tape is walked into the CUDA device through the
            C ABI (`execute_tape`).  The reference's acceptance run (`pytest --backend=...`,
            frontend/test/conftest.py:39-46) needs PennyLane + JAX + the MLIR toolchain: not runnable on
            this image."""
            from . import capi
            single = not isinstance(circuits, (list, tuple))
            tapes = [circuits] if single else list(circuits)

            def make(n, shots):
                d = capi.B200Device(n, **self.device_kwargs)
                if shots:
                    d.set_shots(shots)
                return d
            res = tuple(execute_tape(t, make, wires=self.wires) for t in tapes)
            return res[0] if single else res
except ImportError:  # PennyLane absent: the C-ABI / runtime-C-API paths still work
    B200StateVectorDevice = None
