"""CPU tests: the oracle's C port of the residual path against (a) the compiled reference when it is
present (oracle/_ref, built from /root/reference) and (b) the committed golden digests generated from it.
Shapes follow the reference's own DSP test, tests/checkasm/vvc_itx.c:58-92."""
import os

import numpy as np
import pytest

from ffmpeg_b200 import synth
from oracle import portapi, refapi
from tests import util

needs_ref = pytest.mark.skipif(not refapi.available(), reason="oracle/_ref not built (no /root/reference on this machine)")


def test_port_itx_matches_golden():
    g = np.load(os.path.join(util.GOLDEN, "itx.npz"))
    rng = np.random.default_rng(int(g["seed"]))
    n = 0
    # replay the generator's random stream exactly (tools/gen_golden.py:golden_itx)
    populated = {tuple(k[:4]) for k in g["keys"]}
    digest = {tuple(k[:4]): d for k, d in zip(g["keys"], g["md5"])}
    for trh in range(3):
        for trv in range(3):
            for l2w in range(7):
                for l2h in range(7):
                    w, h = 1 << l2w, 1 << l2h
                    nzw = int(rng.integers(1, min(w, 32 if trh == 0 else 16) + 1))
                    nzh = int(rng.integers(1, min(h, 32 if trv == 0 else 16) + 1))
                    c = np.zeros((h, w), np.int32)
                    c[:nzh, :nzw] = rng.integers(-32768, 32768, (nzh, nzw))
                    if (trh, trv, l2w, l2h) not in populated:
                        continue
                    flat = c.reshape(-1).copy()
                    portapi.itx(flat, trh, trv, l2w, l2h, nzw, nzh)
                    assert util.md5(flat) == digest[(trh, trv, l2w, l2h)], (trh, trv, l2w, l2h, nzw, nzh)
                    n += 1
    assert n == 210      # 210 populated itx slots


def test_port_residual_matches_golden():
    g = np.load(os.path.join(util.GOLDEN, "residual.npz"))
    for name in ("cif_realistic", "wqvga_full"):
        seed, w, h, full = (int(v) for v in g[name + "_args"])
        pic = synth.residual_picture(seed, w, h, levels="full" if full else "realistic")
        planes = util.noise_planes(seed, pic["dims"])
        res = portapi.residual_batch(planes, pic["levels"], pic["tbs"], pic["params"], True)
        assert util.plane_md5(planes) == list(g[name + "_plane_md5"])
        assert util.md5(res) == str(g[name + "_residual_md5"])


@needs_ref
@pytest.mark.parametrize("seed", range(4))
def test_port_residual_matches_reference(seed):
    pic = synth.residual_picture(100 + seed, 208, 120, levels="full" if seed & 1 else "realistic")
    a = util.noise_planes(seed, pic["dims"])
    b = [p.copy() for p in a]
    ra = refapi.residual_batch(a, pic["levels"], pic["tbs"], pic["params"], pic["intra_mode"], True)
    rb = portapi.residual_batch(b, pic["levels"], pic["tbs"], pic["params"], True)
    assert np.array_equal(ra, rb)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


@needs_ref
def test_port_itx_all_slots_random_nz():
    """All 210 slots x random (nzw, nzh), values clipped to log2_transform_range like checkasm (vvc_itx.c:68-76)."""
    rng = np.random.default_rng(7)
    n = 0
    for trh in range(3):
        for trv in range(3):
            for l2w in range(7):
                for l2h in range(7):
                    w, h = 1 << l2w, 1 << l2h
                    probe = np.zeros(w * h, np.int32)
                    if not refapi.itx(probe, trh, trv, l2w, l2h, 1, 1):
                        continue
                    n += 1
                    for _ in range(3):
                        nzw = int(rng.integers(1, min(w, 32 if trh == 0 else 16) + 1))
                        nzh = int(rng.integers(1, min(h, 32 if trv == 0 else 16) + 1))
                        c = np.zeros((h, w), np.int32)
                        c[:nzh, :nzw] = rng.integers(-32768, 32768, (nzh, nzw))
                        a, b = c.reshape(-1).copy(), c.reshape(-1).copy()
                        refapi.itx(a, trh, trv, l2w, l2h, nzw, nzh)
                        portapi.itx(b, trh, trv, l2w, l2h, nzw, nzh)
                        assert np.array_equal(a, b), (trh, trv, l2w, l2h, nzw, nzh)
    assert n == 210


@needs_ref
def test_golden_fixtures_are_current():
    """The committed fixtures are exactly what the compiled reference produces today."""
    import subprocess
    import sys
    root = os.path.dirname(util.GOLDEN.rstrip("/")).rsplit("/tests", 1)[0]
    subprocess.check_call([sys.executable, os.path.join(root, "tools", "gen_golden.py"), "--check"], cwd=root)
