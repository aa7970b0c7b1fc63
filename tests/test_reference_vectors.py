"""The slot for vectors from the real Hivemall jar (tests/golden/hivemall/README.md): one test per
clause of the arithmetic contract, against the CPU oracle (no GPU needed) and against the CUDA path
(-m gpu).  Skipped with "no reference vectors supplied" while the slot is empty -- which it is:
PARITY IS UNPINNED.  The kit itself is exercised on templates generated from
the oracle, so that the day real vectors arrive the only open question is the numbers."""
import os
import subprocess
import sys

import pytest

import _golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KINDS = sorted(_golden.CHECKERS)


def _run(kind, impl, directory=_golden.SLOT, allow_templates=False):
    found = _golden.files(kind, directory, allow_templates)
    if not found:
        pytest.skip("no reference vectors supplied")
    for path, j in found:
        try:
            _golden.CHECKERS[kind](j, impl)
        except AssertionError as e:
            raise AssertionError(f"{os.path.basename(path)}: {e}") from None
    return len(found)


@pytest.mark.parametrize("kind", KINDS)
def test_oracle_against_reference_vectors(kind):
    _run(kind, _golden.OracleImpl())


@pytest.mark.gpu
@pytest.mark.parametrize("kind", KINDS)
def test_cuda_path_against_reference_vectors(kind):
    _run(kind, _golden.GpuImpl())


@pytest.fixture(scope="module")
def templates(tmp_path_factory):
    d = str(tmp_path_factory.mktemp("golden_templates"))
    subprocess.check_call([sys.executable, os.path.join(ROOT, "scripts", "golden_template.py"), d])
    return d


@pytest.mark.parametrize("kind", KINDS)
def test_kit_on_oracle_templates(kind, templates):
    """The loader and checkers work (oracle vs files the oracle wrote), and marked templates are ignored
    by the real tests."""
    assert _run(kind, _golden.OracleImpl(), templates, allow_templates=True) >= 1
    assert _golden.files(kind, templates, allow_templates=False) == []


@pytest.mark.gpu
@pytest.mark.parametrize("kind", KINDS)
def test_kit_on_oracle_templates_cuda(kind, templates):
    """The same files through the C-ABI: what will run the day real vectors arrive."""
    assert _run(kind, _golden.GpuImpl(), templates, allow_templates=True) >= 1
