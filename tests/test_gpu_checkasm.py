"""the checkasm-style view of the boundary.  oracle/shimlib/dsp_shim.c fills a VVCDSPContext whose
members push one block through libvvcb200.so (vvcb200_dsp_init, the mirror of ff_vvc_dsp_init, vvcdsp.h:170), and its
driver — modelled on tests/checkasm/vvc_itx.c:58-92 and vvc_mc.c:68-324 — calls the reference's table and that one with
the same random inputs over the reference tests' own shapes."""
import pytest

from oracle import refapi

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not refapi.shim_available(), reason="oracle/_ref/libvvcshim.so not built")]


@pytest.mark.parametrize("bd", [10, 12])
def test_every_member_with_a_device_view_matches_the_reference_table(bd):
    res, device_errors, unrepresentable = refapi.shim_check(bd, seed=bd)
    assert device_errors == 0 and unrepresentable == 0
    members = {f: m for f, (m, _, _) in res.items()}
    # 210 transform members (vvcdsp.c:94-195); of each put flavour's 2 x 7 x 2 x 2 members the 2 x 6 x 2 x 2 that a block
    # width can select (luma 4..128, chroma 2..64, as tests/checkasm/vvc_mc.c sweeps them); the two single members
    assert members == {"itx.itx": 210, "itx.add_residual": 1, "lmcs.filter": 1, "inter.put": 48, "inter.put_uni": 48,
                       "inter.put_uni_w": 48}
    for family, (_, calls, bad) in res.items():
        assert calls > 0 and bad == 0, "%s: %d of %d calls differ" % (family, bad, calls)
