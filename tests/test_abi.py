"""CPU tests of the drop-in boundary: the shared library loads without a GPU and exports every symbol that
include/vvc_b200.h declares; POD struct sizes match between C and the Python mirrors."""
import ctypes as C
import os
import re
import subprocess
import tempfile

from ffmpeg_b200 import abi, lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "vvc_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vvcb200_\w+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    l = lib.load()
    syms = declared_symbols()
    assert len(syms) >= 6
    for s in syms:
        assert hasattr(l, s), "libvvcb200.so does not export %s" % s


def test_abi_version_and_error_string():
    l = lib.load()
    assert l.vvcb200_abi_version() == 2
    assert isinstance(l.vvcb200_last_error(), bytes)


def test_struct_sizes_match_c():
    names = ["VVCB200TransformBlock", "VVCB200ResidualParams", "VVCB200Planes", "VVCB200PredBlock", "VVCB200InterParams",
             "VVCB200DeblockParams", "VVCB200DeblockTables", "VVCB200SaoCtb", "VVCB200AlfCtb", "VVCB200AlfParams",
             "VVCB200IntraBlock", "VVCB200IntraParams", "VVCB200ProfTable"]
    code = '#include <stdio.h>\n#include "vvc_b200.h"\nint main(void){' + \
        "".join('printf("%%zu\\n", sizeof(%s));' % n for n in names) + "return 0;}\n"
    with tempfile.TemporaryDirectory() as d:
        src = os.path.join(d, "s.c")
        open(src, "w").write(code)
        exe = os.path.join(d, "s")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), src, "-o", exe])
        sizes = [int(v) for v in subprocess.check_output([exe]).split()]
    assert sizes == [abi.TB_DTYPE.itemsize, C.sizeof(abi.ResidualParams), C.sizeof(abi.Planes), abi.PB_DTYPE.itemsize,
                     C.sizeof(abi.InterParams), C.sizeof(abi.DeblockParams), C.sizeof(abi.DeblockTables),
                     abi.SAO_DTYPE.itemsize, abi.ALF_CTB_DTYPE.itemsize, abi.ALF_PARAMS_DTYPE.itemsize,
                     abi.IB_DTYPE.itemsize, C.sizeof(abi.IntraParams), abi.PROF_DTYPE.itemsize]


def test_bad_arguments_are_rejected_without_a_gpu():
    l = lib.load()
    p = abi.residual_params(bit_depth=9)
    rc = l.vvcb200_residual_batch(None, None, None, None, 0, C.byref(p), None)
    assert rc == -1
    pl = abi.Planes()
    rc = l.vvcb200_residual_batch(C.byref(pl), None, None, None, 0, C.byref(p), None)
    assert rc == -1 and b"bit depth" in l.vvcb200_last_error()


def test_sparse_level_packer_round_trips():
    """vvcb200_pack_levels_sparse: bitmap + non-zero values per TB decode back to the dense scan boxes."""
    import numpy as np
    from ffmpeg_b200 import synth
    pic = synth.residual_picture(3, 416, 240)
    pic["levels"][::3] = 0
    lv, tbs = lib.pack_levels_sparse(pic["levels"], pic["tbs"])
    assert (tbs["coeff_off"] % 2 == 0).all()
    for t_old, t_new in list(zip(pic["tbs"], tbs))[::37]:
        n = int(t_old["nzw"]) * int(t_old["nzh"])
        dense = pic["levels"][t_old["coeff_off"]:t_old["coeff_off"] + n]
        words = (n + 31) // 32
        off = int(t_new["coeff_off"])
        bitmap = lv[off:off + 2 * words].view(np.uint32)
        bits = np.unpackbits(bitmap.view(np.uint8), bitorder="little")[:n].astype(bool)
        assert np.array_equal(bits, dense != 0)
        nz = lv[off + 2 * words:off + 2 * words + int(bits.sum())]
        assert np.array_equal(nz, dense[bits])


def test_narrow_level_packer_round_trips():
    """vvcb200_pack_levels with narrow levels: blocks whose levels fit int8 are stored as int8 (flagged TB_LEVELS8), the
    others as int16; both the dense and the sparse form decode back to the scan boxes."""
    import numpy as np
    from ffmpeg_b200 import abi, synth
    pic = synth.residual_picture(5, 416, 240)
    pic["levels"][1::4] = 0
    for sparse in (False, True):
        lv, tbs = lib.pack_levels(pic["levels"], pic["tbs"], sparse=sparse, narrow=True)
        narrow = (tbs["flags"] & abi.TB_LEVELS8) != 0
        assert narrow.any() and not narrow.all()                 # the synthetic levels reach +-200
        assert lv.nbytes < 0.8 * (pic["levels"].nbytes if not sparse else lib.pack_levels_sparse(pic["levels"], pic["tbs"])[0].nbytes)
        if sparse:
            assert (tbs["coeff_off"] % 2 == 0).all()
        for t_old, t_new in list(zip(pic["tbs"], tbs))[::23]:
            n = int(t_old["nzw"]) * int(t_old["nzh"])
            dense = pic["levels"][t_old["coeff_off"]:t_old["coeff_off"] + n]
            is8 = bool(t_new["flags"] & abi.TB_LEVELS8)
            assert is8 == bool(np.all((dense >= -128) & (dense <= 127)))
            off = int(t_new["coeff_off"])
            if sparse:
                words = (n + 31) // 32
                bits = np.unpackbits(lv[off:off + 2 * words].view(np.uint8), bitorder="little")[:n].astype(bool)
                assert np.array_equal(bits, dense != 0)
                want, off = dense[bits], off + 2 * words
            else:
                want = dense
            got = lv[off:off + (len(want) + 1) // 2].view(np.int8)[:len(want)] if is8 else lv[off:off + len(want)]
            assert np.array_equal(got.astype(np.int16), want)


def test_checkasm_shim_loads_and_says_what_it_leaves_to_the_reference():
    """oracle/_ref/libvvcshim.so: exports vvcb200_dsp_init / vvcshim_check and lists, with reasons, the
    VVCDSPContext members that have no block-level device view (no compute here)."""
    import pytest
    from oracle import refapi
    if not refapi.shim_available():
        pytest.skip("oracle/_ref/libvvcshim.so not built")
    s = refapi.shim()
    assert hasattr(s, "vvcb200_dsp_init") and hasattr(s, "vvcshim_check")
    text = s.vvcshim_report().decode()
    for member in ("inter.avg", "sao.band_filter", "alf.filter", "lf.filter_luma", "intra.", "inter.sad"):
        assert member in text
