import os
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, REPO)
sys.path.insert(0, HERE)

ORACLE_LIB = os.path.join(REPO, "oracle", "libartis_oracle.so")
GPU_LIB = os.path.join(REPO, "artis_b200", "libartis_gpu.so")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle_lib():
    """oracle/libartis_oracle.so -- the CPU restatement (test infrastructure). Built on demand (g++ only)."""
    if not os.path.exists(ORACLE_LIB):
        subprocess.run(["make", "-C", os.path.join(REPO, "oracle"), "restatement"], check=True)
    return ORACLE_LIB


@pytest.fixture(scope="session")
def gpu_lib():
    if not os.path.exists(GPU_LIB):
        pytest.fail(f"{GPU_LIB} is missing: run __graft_entry__.build() (the product has no CPU fallback)")
    return GPU_LIB


@pytest.fixture(scope="session", params=["classic_nogrey_ts12.npz", "classic_grey_ts8.npz", "kilonova_lte_ts1.npz",
                                        "nebular_nlte_ts8.npz", "classic_1d_ts12.npz", "classic_vpkt_ts30.npz"])
def fixture_name(request):
    return request.param
