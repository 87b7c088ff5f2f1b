"""CPU-side checks of the plugin boundary: the reference's runtime C API builds from its sources,
drives a conforming plugin (the reference's null.qubit) through tests/rt_driver.py, and our plugin
exports the factory symbol the runtime will dlsym (reference: ExecutionContext.hpp:254-267)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import rt_driver

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    from catalyst_b200 import build
    from oracle import build_ref
    build.build_all()
    build_ref.build()
    if not (os.path.exists(rt_driver.RT_CAPI) and os.path.exists(rt_driver.PLUGIN)):
        pytest.skip("reference sources not available and no prebuilt librt_capi / plugin")


def test_plugin_exports_factory(built):
    out = subprocess.check_output(["nm", "-D", "--defined-only", rt_driver.PLUGIN]).decode()
    assert " T B200StateVectorFactory" in out
    # the adapter resolves every b200sv_* it uses from libb200sv.so next to it (rpath $ORIGIN)
    L = C.CDLL(rt_driver.PLUGIN)
    assert hasattr(L, "B200StateVectorFactory")


def test_driver_against_reference_null_qubit(built):
    # same driver, the reference's own conforming plugin: validates the harness, not our device
    d = rt_driver.RuntimeDevice(3, lib=rt_driver.NULL_PLUGIN, name="NullQubit")
    try:
        assert d.num_qubits() == 3
        d.named_op("Hadamard", [], [0])
        d.named_op("CRX", [0.3], [0, 1], inverse=True, ctrl_wires=[2], ctrl_values=[True])
        d.named_op("MultiRZ", [0.3], [0, 1, 2])
        k = d.obs_hamiltonian([0.5], [d.obs_tensor([d.obs_named("PauliZ", 0), d.obs_named("PauliX", 1)])])
        assert d.expval(k) == 0.0
        p = d.probs()
        assert p[0] == 1.0 and p.shape == (8,)
        assert d.measure(0) in (False, True)
        # dynamic qubits through __catalyst__rt__qubit_allocate / qubit_release and a second QirArray
        # (RuntimeCAPI.cpp:305-372): the calls tests/test_gpu_zz_reference_programs.py makes on our plugin
        d.allocate_qubit()
        assert d.num_qubits() == 4 and len(d.q) == 4
        d.named_op("CNOT", [], [3, 0])
        d.release_qubit(3)
        assert d.num_qubits() == 3 and len(d.q) == 3
        reg, w = d.allocate_register(2)
        assert w == [3, 4] and d.num_qubits() == 5
        d.named_op("CNOT", [], [w[1], 0])
        d.release_register(reg, w)
        assert d.num_qubits() == 3 and len(d.q) == 3
    finally:
        d.close()


def test_toml_capabilities_file():
    import tomllib  # noqa: F401  (py3.11+)
    path = os.path.join(ROOT, "catalyst_b200", "b200sv.toml")
    cfg = tomllib.load(open(path, "rb"))
    assert cfg["schema"] == 3
    gates = cfg["operators"]["gates"]
    for g in ["PauliX", "Hadamard", "RX", "CNOT", "Toffoli", "MultiRZ", "QubitUnitary", "PauliRot",
              "DoubleExcitation", "PCPhase", "GlobalPhase", "ControlledPhaseShift"]:
        assert g in gates, g
    assert "differentiable" not in gates["Rot"].get("properties", [])
    assert cfg["compilation"]["qjit_compatible"] is True
