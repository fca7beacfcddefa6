"""CPU: the built library's SASS holds the Blackwell instructions the design claims (cuobjdump of libpcc_b200.so, no GPU
needed): tcgen05.mma (UTCHMMA), TMEM loads (LDTM), tcgen05.commit (UTCBAR) and bulk copies (UBLKCP) in both convolution
kernels, cp.async gathers (LDGSTS), no legacy mma.sync (HMMA) anywhere, and the committed opcode listing is current."""
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "demo-learned-point-cloud-compression_b200", "b200pcc", "libpcc_b200.so")

pytestmark = pytest.mark.skipif(shutil.which("cuobjdump") is None or not os.path.exists(SO),
                                reason="needs cuobjdump and the built library")


@pytest.fixture(scope="module")
def table():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sass_opcodes.py")], capture_output=True, text=True,
                         check=True).stdout.splitlines()
    head = out[1].split()
    cols = head[2:]                      # after 'kernel', 'insts'
    rows = {}
    for line in out[2:]:
        parts = line.rsplit(None, len(cols) + 1)
        if len(parts) == len(cols) + 2:
            rows[parts[0].strip()] = dict(zip(["insts"] + cols, map(int, parts[1:])))
    return rows


def test_conv_kernels_use_tcgen05_tmem_and_bulk_copies(table):
    conv = {k: v for k, v in table.items() if "k_spconv_octet<" in k or "k_spconv_umma<" in k}
    assert len(conv) >= 40
    for name, c in conv.items():
        assert c["UTCHMMA"] >= 1, name          # tcgen05.mma
        assert c["LDTM"] >= 1, name             # tcgen05.ld (TMEM -> registers in the epilogue)
        assert c["UTCBAR"] >= 1, name           # tcgen05.commit -> mbarrier
        assert c["UBLKCP"] >= 1, name           # cp.async.bulk (weight panels / kernel-map ring)
        assert c["LDGSTS"] >= 1, name           # cp.async row gather
        assert c["SYNCS"] >= 1, name            # mbarrier hand-offs
    octet = [c for k, c in conv.items() if "k_spconv_octet<" in k]
    assert all(c["UTCHMMA"] >= 288 for c in octet)      # the compile-time MMA schedule of a tile (288 + 16 residual)


def test_no_legacy_tensor_core_path_and_listing_is_current(table):
    assert "HMMA" not in next(iter(table.values())) or all(c.get("HMMA", 0) == 0 for c in table.values())
    total = table["TOTAL"]
    assert total["UTCHMMA"] > 1000 and total["VOTE"] > 100 and total["SHFL"] > 100      # warp primitives in the coder / top-k
    listed = open(os.path.join(ROOT, "profiles", "sass_opcodes.txt")).read()
    for name in ("pcc::k_spconv_octet<", "pcc::k_spconv_umma<", "pcc::k_rans_decode<", "pcc::k_kmap_k3_blocks"):
        assert name in listed
