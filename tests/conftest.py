"""Shared fixtures.  GPU tests are marked `gpu`; everything else runs on a CPU-only machine.

The CPU oracle (oracle/_build/liboracle.so) is the checker; the product is gffx_b200 (ctypes over
gffx_b200/lib/libgffx_b200.so).  Nothing here reads /root/reference at run time — the reference's
real index files used as fixtures are committed under tests/golden/ (see tests/golden/make_golden.py).
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


def _ensure_built():
    lib = os.path.join(ROOT, "gffx_b200", "lib", "libgffx_b200.so")
    orc = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
    if not (os.path.exists(lib) and os.path.exists(orc)):
        subprocess.check_call(["make", "-j8"], cwd=ROOT, stdout=subprocess.DEVNULL)


_ensure_built()


@pytest.fixture(scope="session")
def oracle():
    from tests import _oracle

    return _oracle


@pytest.fixture(scope="session")
def gffx():
    import gffx_b200

    return gffx_b200


def _make_synth(tmp, name, seed, n_genes, quirks):
    import gffx_b200
    from tests import _oracle

    path = os.path.join(str(tmp), name)
    gffx_b200.synth_gff(path, seed=seed, n_genes=n_genes, quirks=quirks)
    _oracle.build_index(path)
    return path


@pytest.fixture(scope="session")
def synth_quirk_gff(tmp_path_factory):
    """~90k-line human-like GFF3 with every quirk of SURVEY.md App. D switched on + its index files."""
    return _make_synth(tmp_path_factory.mktemp("quirk"), "quirk.gff", 7, 2000, True)


@pytest.fixture(scope="session")
def synth_plain_gff(tmp_path_factory):
    return _make_synth(tmp_path_factory.mktemp("plain"), "plain.gff", 11, 3000, False)


@pytest.fixture(scope="session")
def kat_gff(tmp_path_factory):
    """The hand-derived known-answer fixture of SURVEY.md App. C."""
    d = tmp_path_factory.mktemp("kat")
    path = os.path.join(str(d), "kat.gff")
    with open(os.path.join(GOLDEN, "kat.gff"), "rb") as f, open(path, "wb") as g:
        g.write(f.read())
    from tests import _oracle

    _oracle.build_index(path)
    return path


@pytest.fixture(scope="session")
def ps_index(tmp_path_factory):
    """The reference's real `benchmark/ps.gff.{sqs,rix,rit,gof}` (no GFF text: it is a git-LFS pointer upstream)."""
    import lzma
    import shutil

    d = tmp_path_factory.mktemp("ps")
    for ext in ("sqs", "rix", "rit", "gof"):
        shutil.copy(os.path.join(GOLDEN, "ps", "ps.gff." + ext), os.path.join(str(d), "ps.gff." + ext))
    bed = os.path.join(str(d), "input_001.bed")
    with lzma.open(os.path.join(GOLDEN, "ps", "input_001.bed.xz"), "rb") as f, open(bed, "wb") as g:
        g.write(f.read())
    return os.path.join(str(d), "ps.gff"), bed


def has_cuda():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False
