import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


# The parity tests target the large-request pipelines (k_path / k_primary + the wavefront kernels), the ones the headline numbers
# come from; requests of the tests' size would otherwise run as ONE k_mega launch (bf_set_tunable "mega_max").  k_mega has its own
# tests: bit-identical records against the wavefront sequence (test_mega_kernel_equals_wavefront) and the oracle comparison of
# test_profile_small / test_sample_records_small with pipeline = "mega".  Read once per scene, in bf_scene_create.
os.environ.setdefault("BF_MEGA_MAX", "-1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def native():
    """The built native library; building is the job of __graft_entry__.build()."""
    from beifong_b200 import _native
    if not os.path.exists(_native.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return _native.lib()
