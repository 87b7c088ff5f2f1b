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
                     world_size=1, rank=0, comm=None):
            super().__init__(wires=wires, shots=shots)
            self.device_kwargs = {"fuse": bool(fuse)}
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
            raise NotImplementedError(
                "b200.statevector executes through Catalyst's runtime (@qjit); interpreted "
                "execution is a 'next' row (SURVEY.md 8(f) N1)")
except ImportError:  # PennyLane absent: the C-ABI / runtime-C-API paths still work
    B200StateVectorDevice = None
