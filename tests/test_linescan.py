"""the GFF line table built by the device scanner (gffx_b200/csrc/linescan.cu) and read back from the
side-car cache (`<gff>.b200cache`, linecache.cpp) must be the table the host scanner builds: same digest over every column,
the three dictionaries and the per-block facts — on the quirk fixture (CRLF lines, lines with < 9 columns, end == 0, a
block that is not valid UTF-8, duplicate IDs) as on plain and scaffold-rich texts."""
import os
import shutil

import numpy as np
import pytest

from tests._oracle import Oracle


class _Env:
    def __init__(self, **kv):
        self.kv = kv

    def __enter__(self):
        self.old = {k: os.environ.get(k) for k in self.kv}
        for k, v in self.kv.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = str(v)

    def __exit__(self, *a):
        for k, v in self.old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def _info(gffx, path, device, **env):
    with _Env(**env):
        idx = gffx.TreeIndexData.load_tree_index(path, device=device)
    try:
        return idx.lines_info() + (idx.n_lines, idx.n_ids)
    finally:
        idx.close()


def test_cache_roundtrip_on_the_host(gffx, synth_quirk_gff, tmp_path):
    d = str(tmp_path)
    cold = _info(gffx, synth_quirk_gff, -1, GFFX_CACHE_DIR=d, GFFX_CACHE=None)
    assert cold[1] == 0 and os.path.exists(os.path.join(d, os.path.basename(synth_quirk_gff) + ".b200cache"))
    warm = _info(gffx, synth_quirk_gff, -1, GFFX_CACHE_DIR=d, GFFX_CACHE=None)
    assert warm[1] == 2 and warm[0] == cold[0] and warm[2:] == cold[2:]
    off = _info(gffx, synth_quirk_gff, -1, GFFX_CACHE_DIR=d, GFFX_CACHE=0)
    assert off[1] == 0 and off[0] == cold[0]


def test_stale_or_damaged_cache_is_ignored(gffx, synth_plain_gff, tmp_path):
    """the key is the GFF's size + mtime and the block list: another text, a truncated or a scribbled cache file = no cache"""
    d = str(tmp_path / "c")
    os.makedirs(d)
    g = str(tmp_path / "p.gff")
    for ext in ("", ".sqs", ".rix", ".rit", ".gof"):
        shutil.copy(synth_plain_gff + ext, g + ext)
    a = _info(gffx, g, -1, GFFX_CACHE_DIR=d)
    cache = os.path.join(d, "p.gff.b200cache")
    assert _info(gffx, g, -1, GFFX_CACHE_DIR=d)[1] == 2
    # same bytes, newer mtime
    st = os.stat(g)
    os.utime(g, ns=(st.st_atime_ns, st.st_mtime_ns + 10**9))
    b = _info(gffx, g, -1, GFFX_CACHE_DIR=d)
    assert b[1] == 0 and b[0] == a[0]
    assert _info(gffx, g, -1, GFFX_CACHE_DIR=d)[1] == 2        # ... and it was rewritten for the new key
    # truncated / damaged files
    blob = open(cache, "rb").read()
    open(cache, "wb").write(blob[: len(blob) // 2])
    assert _info(gffx, g, -1, GFFX_CACHE_DIR=d)[:2] == (a[0], 0)
    open(cache, "wb").write(b"GFXB200C" + blob[8:40][::-1] + blob[40:])
    assert _info(gffx, g, -1, GFFX_CACHE_DIR=d)[:2] == (a[0], 0)
    # an unwritable cache directory is not an error
    assert _info(gffx, g, -1, GFFX_CACHE_DIR=str(tmp_path / "does" / "not" / "exist"))[:2] == (a[0], 0)


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["quirk", "plain", "scaffolds"])
def test_device_scanner_builds_the_host_scanners_table(gffx, synth_quirk_gff, synth_plain_gff, tmp_path, which):
    if which == "scaffolds":
        from tests import _oracle
        path = str(tmp_path / "s.gff")
        gffx.synth_gff(path, seed=3, n_genes=1500, quirks=True, n_scaffolds=700)
        _oracle.build_index(path)
    else:
        path = synth_quirk_gff if which == "quirk" else synth_plain_gff
    host = _info(gffx, path, 0, GFFX_CACHE=0, GFFX_LINE_SCAN="host")
    dev = _info(gffx, path, 0, GFFX_CACHE=0, GFFX_LINE_SCAN="device")
    assert host[1] == 0 and dev[1] == 1
    assert dev[0] == host[0] and dev[2:] == host[2:]
    # the device scan with the ID dictionary built by the host instead of the device kernels (the path a hash collision takes)
    dev_h = _info(gffx, path, 0, GFFX_CACHE=0, GFFX_LINE_SCAN="device", GFFX_ID_DICT="host")
    assert dev_h[1] == 1 and dev_h[0] == host[0] and dev_h[2:] == host[2:]
    # cold device scan -> cache -> warm open: the same table a third time
    d = str(tmp_path / "cache")
    os.makedirs(d)
    cold = _info(gffx, path, 0, GFFX_CACHE_DIR=d, GFFX_LINE_SCAN="device")
    warm = _info(gffx, path, 0, GFFX_CACHE_DIR=d, GFFX_LINE_SCAN="device")
    assert (cold[1], warm[1]) == (1, 2) and cold[0] == warm[0] == host[0]


@pytest.mark.gpu
def test_results_after_a_device_scan(gffx, synth_quirk_gff):
    """the engine adopts the text the scanner left in HBM: intersect, record emission, depth and coverage as from a host-scanned index"""
    with _Env(GFFX_CACHE=0, GFFX_LINE_SCAN="device"):
        idx = gffx.TreeIndexData.load_tree_index(synth_quirk_gff, device=0)
    o = Oracle(synth_quirk_gff)
    try:
        assert idx.lines_info()[1] == 1
        regs = idx.synth_regions(11, 4000, 20000)
        for mode, invert, entire, types in [(2, False, False, None), (0, True, False, "gene,mRNA,exon"), (1, False, True, None)]:
            r = idx.intersect(regs, mode, invert, entire, types)
            want, n_pairs, _ = o.intersect(regs, mode, invert, entire, types)
            assert r.n_pairs == n_pairs and np.array_equal(r.ranges, want)
        r = idx.intersect(regs, 2, False, False, None)
        text = idx.gather_text(r.ranges)
        gff = open(synth_quirk_gff, "rb").read()
        assert bytes(text) == b"".join(gff[int(a):int(b)] for a, b in r.ranges)
        reads = idx.synth_regions(12, 30000, 150)
        assert idx.depth(reads).to_tsv() == o.depth(reads)
        assert idx.coverage(reads).to_tsv() == o.coverage(reads)
    finally:
        o.close()
        idx.close()
