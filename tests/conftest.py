import os
import sys
import warnings

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

warnings.filterwarnings("ignore", category=DeprecationWarning, message=".*audioop.*")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA GPU (B200); run with -m gpu on the GPU box")
    # A fresh checkout has no built library (build outputs are not tracked): build it once, as
    # __graft_entry__.build() does (nvcc cross-compiles sm_100a without a GPU).  The product
    # itself never does this: importing it without the library fails loudly.
    lib = os.path.join(ROOT, "synthesizer_b200", "libsynth_b200.so")
    if not os.path.exists(lib):
        import __graft_entry__
        __graft_entry__.build()


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("a test marked gpu ran without a CUDA device")
    return torch.device("cuda", 0)
