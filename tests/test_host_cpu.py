"""Host-side logic that needs no GPU: the C-ABI library loads and exports every symbol
include/lmc_b200.h declares, struct layouts agree with the header, filter padding / tap
extraction, roofline arithmetic, error behaviour without a device."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header():
    with open(os.path.join(ROOT, "include", "lmc_b200.h")) as f:
        return f.read()


def test_library_exports_every_declared_symbol():
    from lmc_atomi_b200 import _lib, build
    build.build()
    lib = _lib.load()
    src = re.sub(r"/\*.*?\*/", "", _header(), flags=re.S)
    names = set(re.findall(r"\b(lmc_[a-z0-9_]+)\s*\(", src))
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/lmc_b200.h but not exported"
    assert set(_lib.SIGNATURES) <= names | {"lmc_last_error", "lmc_version"}
    assert lib.lmc_version() >= 100


def test_ctypes_struct_sizes_match_header_layout():
    from lmc_atomi_b200 import _lib
    # sizes computed by hand from the header (natural alignment, LMC_MAX_COMP=8, LMC_IMG_MAX_TAPS=16)
    assert ctypes.sizeof(_lib.MixTarget) == 8 + 8 * (2 + 4 + 1 + 1 + 1) * 8 + 8 + 8 + 16
    assert ctypes.sizeof(_lib.MixStep) == 8 + 8 + 4 * 32 + 8 + 16 + 16 + 8
    assert ctypes.sizeof(_lib.ImgProblem) == 24 + 2 * 16 * 8 + 8 + 8 + 8
    assert ctypes.sizeof(_lib.ImgMyulaParams) == 5 * 8 + 8
    assert ctypes.sizeof(_lib.ImgUlpdaParams) == 4 * 8 + 4 * 4 + 3 * 8


def test_no_device_is_a_loud_error():
    import torch
    if torch.cuda.is_available():
        pytest.skip("has a device")
    from lmc_atomi_b200 import _lib
    with pytest.raises(_lib.LmcError):
        _lib.require_cuda()
    from lmc_atomi_b200.lmc import LangevinMonteCarlo
    s = LangevinMonteCarlo([np.zeros(2)], [np.eye(2)], [1.0], K=10)
    with pytest.raises(_lib.LmcError):
        s.ula(0.1)


@pytest.mark.parametrize("k", [5, 6, 7])
def test_pad_filter_and_taps(k):
    from lmc_atomi_b200 import imaging
    from oracle import deconv_ref as D
    h = np.ones((k, k)) / k ** 2
    hp = imaging.pad_filter(h, (k // 2, k // 2))
    assert np.array_equal(hp, D.pad_kernel_like_pylops(h, (k // 2, k // 2)))
    assert hp.shape[0] % 2 == 1 and hp.shape[1] % 2 == 1
    wy, lo_y, wx, lo_x = imaging.separable_taps(hp)
    assert len(wy) == len(wx) == k and np.allclose(np.outer(wy, wx), h)
    assert (lo_y, lo_x) == {5: (-2, -2), 6: (-2, -2), 7: (-3, -3)}[k]
    # the taps reproduce scipy's 'same' convolution on an impulse
    dims = (15, 17)
    x = np.zeros(dims)
    x[7, 8] = 1.0
    want = D.Convolve2D(dims, h, offset=(k // 2, k // 2)).matvec(x.ravel()).reshape(dims)
    got = np.zeros(dims)
    for i in range(dims[0]):
        for j in range(dims[1]):
            for a, wa in enumerate(wy):
                for b, wb in enumerate(wx):
                    ii, jj = i + lo_y + a, j + lo_x + b
                    if 0 <= ii < dims[0] and 0 <= jj < dims[1]:
                        got[i, j] += wa * wb * x[ii, jj]
    assert np.abs(got - want).max() < 1e-15


def test_non_separable_filter_is_refused():
    from lmc_atomi_b200 import imaging
    h = np.array([[1.0, 0, 0], [0, 1, 0], [0, 0, 1]]) / 3
    with pytest.raises(NotImplementedError):
        imaging.separable_taps(h)


def test_roofline_counts():
    from lmc_atomi_b200 import roofline as R
    assert R.myula_bytes_per_pixel(10) == 408 and R.ulpda_bytes_per_pixel(9) == 672
    assert R.mix_laneops_per_step("gauss", "ula") == 98 + 58 + 12
    assert abs(R.issue_peak_laneops(1965.0) - 148 * 128 * 1.965e9) < 1
    p = R.measured_peaks()
    assert p["hbm_gbs"] > 1000 and p["source"] in ("measured", "fallback")


def test_product_package_never_imports_the_oracle():
    """The oracle is test infrastructure: no module of the product package may import it."""
    import os
    import re
    pkg = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lmc_atomi_b200")
    bad = []
    for name in sorted(os.listdir(pkg)):
        if name.endswith(".py"):
            src = open(os.path.join(pkg, name)).read()
            if re.search(r"^\s*(from\s+oracle\b|import\s+oracle\b)", src, flags=re.M):
                bad.append(name)
    assert not bad, bad
