"""Shared fixtures.  `-m "not gpu"` = oracle vs reference/golden vectors, host logic, ABI surface;
`-m gpu` = the parity tests proper (CUDA path through the C ABI vs the oracle)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import __graft_entry__ as entry  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    if not os.path.exists(entry.LIB):
        entry.build_library()
    return entry.load_package()


@pytest.fixture(scope="session")
def abi(pkg):
    return pkg.abi


@pytest.fixture(scope="session")
def ob(pkg):
    from oracle import bindings
    return bindings


@pytest.fixture(scope="session")
def port(ob):
    return ob.PortOracle()


@pytest.fixture(scope="session")
def ref(ob):
    if not ob.have_ref():
        pytest.skip("oracle/_ref/libref_oracle.so not built (needs /root/reference)")
    return ob.RefOracle()


@pytest.fixture(scope="session")
def gpu(pkg):
    if pkg.lib().rtow_device_count() < 1:
        pytest.fail("no CUDA device: the product has no CPU path (run with -m 'not gpu' on CPU boxes)")
    return pkg


def earth_image():
    """Stand-in for input_images/1_earth_8k.jpg (a procedural RGB8 image; the reference's JPEG decoder is out of scope)."""
    ny, nx = 64, 128
    y, x = np.mgrid[0:ny, 0:nx]
    img = np.stack([(x * 2) % 256, (y * 4) % 256, (x + y) % 256], axis=-1).astype(np.uint8)
    return img


SCENES = [
    ("random_ned", dict(wrap=True)),
    ("random", dict(wrap=False)),
    ("cornell_ned", dict(wrap=True)),
    ("cornell", dict(wrap=False)),
    ("cornell_smoke_ned", dict(wrap=True)),
]


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
