import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "ref: needs the compiled reference under oracle/_ref (skipped when absent)")


# GPU tests fail loudly (they do not skip) when selected with -m gpu on a box whose GPU or extension is missing; without
# -m gpu they are deselected by the marker expression the driver passes.


@pytest.fixture(scope="session")
def golden():
    import json
    return json.loads((ROOT / "tests" / "golden" / "golden.json").read_text())


@pytest.fixture(scope="session")
def golden_case():
    def load(name: str) -> bytes:
        return (ROOT / "tests" / "golden" / "cases" / f"{name}.bz2").read_bytes()
    return load
