"""Host-side tile generation of the wavelet noise (mantaflow_b200/csrc/noise.cu generateTile, no GPU needed) against the tile the
unmodified reference wrote (tests/golden/noise_tile_ref.npz, made by tests/golden/make_noise_tile_golden.py): bit-exact."""
import ctypes as Ct
import hashlib
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def test_noise_tile_matches_reference():
    from mantaflow_b200._lib import lib, check
    ref = np.load(os.path.join(HERE, "golden", "noise_tile_ref.npz"))
    out = np.empty(3 * 128 ** 3, np.float32)
    check(lib().mf_noise_generate_tile(out.ctypes.data_as(Ct.POINTER(Ct.c_float))))
    assert np.array_equal(out[::1021].view(np.int32), ref["sample"].view(np.int32))
    assert hashlib.sha256(out.tobytes()).hexdigest() == str(ref["sha256"])
    assert out.astype(np.float64).sum() == float(ref["total"])
