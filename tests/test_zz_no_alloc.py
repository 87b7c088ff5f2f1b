"""The allocation-free entry point `b200k_bilinear_pauli_noalloc` (include/b200sv_kernels.h; no counterpart in
the reference): called through the C ABI and compared with the
existing `b200k_bilinear_pauli` on two calls in a row (the second call finds the first call's leftovers in the
shared static scratch).  Collected last on purpose.

Device memory for the tiny output / scratch buffers comes from torch (plumbing only); the state vectors are the
CUDA device's own shards (`b200sv_device_pointer`)."""
import ctypes as C

import numpy as np
import pytest

from dsl import apply_ops, random_circuit

pytestmark = pytest.mark.gpu
TOL = 1e-12


def _bind(L):
    u64, dp = C.c_uint64, C.c_void_p
    L.b200k_bilinear_pauli.argtypes = [dp, dp, C.c_int, u64, u64, u64, C.c_int, u64, u64, dp, dp, dp]
    L.b200k_bilinear_pauli.restype = C.c_int
    L.b200k_bilinear_pauli_noalloc.argtypes = [dp, dp, C.c_int, u64, u64, u64, C.c_int, u64, u64, dp, dp,
                                               C.c_char_p, C.c_int64]
    L.b200k_bilinear_pauli_noalloc.restype = C.c_int
    L.b200k_reduce_scratch_doubles.restype = C.c_int64


def test_noalloc_bilinear_matches_existing_entry_point_twice():
    import torch
    from catalyst_b200 import capi
    L = capi.lib()
    _bind(L)
    n = 16
    da, db = capi.B200Device(n, seed=1, virtual_swap=False), capi.B200Device(n, seed=2, virtual_swap=False)
    apply_ops(da, random_circuit(11, n, 60))
    apply_ops(db, random_circuit(12, n, 60))
    da.synchronize()
    db.synchronize()
    pa, pb = da.device_pointer(), db.device_pointer()
    assert pa and pb
    torch.cuda.set_device(0)
    scratch = torch.zeros(int(L.b200k_reduce_scratch_doubles()), dtype=torch.float64, device="cuda")
    out_ref = torch.zeros(2, dtype=torch.float64, device="cuda")
    out_new = torch.full((2,), 7.0, dtype=torch.float64, device="cuda")
    err = C.create_string_buffer(256)
    # (phi, lam, xmask, zmask, ny, cmask, cval): a diagonal word on one vector, then a general word with a
    # control predicate between two different vectors -- two calls in a row on the shared scratch
    calls = [(pa, pa, 0, 0b1001, 0, 0, 0),
             (pa, pb, (1 << 13) | (1 << 2), (1 << 13) | (1 << 7), 1, 1 << 5, 1 << 5),
             (pb, pb, 1 << 15, 0, 0, 0, 0)]
    for phi, lam, xm, zm, ny, cm, cv in calls:
        rc = L.b200k_bilinear_pauli(phi, lam, n, 0, xm, zm, ny, cm, cv, scratch.data_ptr(), out_ref.data_ptr(), None)
        assert rc == 0
        rc = L.b200k_bilinear_pauli_noalloc(phi, lam, n, 0, xm, zm, ny, cm, cv, out_new.data_ptr(), None, err, 256)
        assert rc == 0, err.value.decode()
        torch.cuda.synchronize()
        want, got = out_ref.cpu().numpy(), out_new.cpu().numpy()
        assert np.max(np.abs(got - want)) <= TOL * max(1.0, float(np.max(np.abs(want)))), (got, want)
    # the device's own Expval agrees with the first word (Z on wires n-1-0 and n-1-3 <-> index bits 0 and 3)
    rc = L.b200k_bilinear_pauli_noalloc(pa, pa, n, 0, 0, 0b1001, 0, 0, 0, out_new.data_ptr(), None, err, 256)
    assert rc == 0
    torch.cuda.synchronize()
    ev = da.expval(da.obs_tensor([da.obs_named("PauliZ", n - 1), da.obs_named("PauliZ", n - 4)]))
    assert abs(out_new.cpu().numpy()[0] - ev) < TOL
    da.close()
    db.close()


def test_noalloc_rejects_sizes_beyond_its_maximum():
    import torch
    from catalyst_b200 import capi
    L = capi.lib()
    _bind(L)
    d = capi.B200Device(4)
    d.synchronize()
    out = torch.zeros(2, dtype=torch.float64, device="cuda")
    err = C.create_string_buffer(256)
    rc = L.b200k_bilinear_pauli_noalloc(d.device_pointer(), d.device_pointer(), 37, 0, 0, 0, 0, 0, 0, out.data_ptr(), None,
                                        err, 256)
    assert rc != 0 and b"n_local outside 0..36" in err.value
    rc = L.b200k_bilinear_pauli_noalloc(d.device_pointer(), d.device_pointer(), 4, 0, 1 << 9, 0, 0, 0, 0, out.data_ptr(),
                                        None, err, 256)
    assert rc != 0 and b"xmask" in err.value
    rc = L.b200k_bilinear_pauli_noalloc(None, d.device_pointer(), 4, 0, 0, 0, 0, 0, 0, out.data_ptr(), None, err, 256)
    assert rc != 0 and b"null" in err.value
    d.close()
