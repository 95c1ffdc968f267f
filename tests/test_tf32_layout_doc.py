"""The MN-major tf32 operand layout written by `conv_wgrad_kernel<4>` (`chunk_addr<true>` in csrc/conv_common.cuh)
against what the tensor core was measured to read (`bench/probe_umma.py` -> profiles/r2/probe_umma_layouts.txt).
CPU-only: re-states the address formula and compares it with the recorded hardware offsets, so a change of the
formula without a new probe run is caught."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RECORD = os.path.join(ROOT, "profiles", "r2", "probe_umma_layouts.txt")
BLOCK = 128 * 128            # kSwBlockBytes: 128 pixel rows x 128 bytes (32 channels)


def chunk_addr_sw(row: int, jj: int) -> int:
    """Byte offset of 16-byte chunk `jj` (4 channels) of pixel row `row`: csrc/conv_common.cuh chunk_addr<true>."""
    sw = (row & 3) << 1
    return row * 128 + (jj >> 3) * BLOCK + (((jj & 7) ^ sw) << 4)


def _section(title):
    rows, on = {}, False
    for line in open(RECORD):
        if line.startswith("=="):
            on = line.strip() == "== " + title
            continue
        m = re.match(r"\s*m=\s*(\d+): byte offsets for k=0\.\.\d+: \[(.*)\]", line)
        if on and m:
            rows[int(m.group(1))] = [int(v) for v in m.group(2).split(",")]
    return rows


@pytest.mark.skipif(not os.path.exists(RECORD), reason="probe record not present")
def test_wgrad_tf32_layout_is_what_the_tensor_core_reads():
    # descriptor of the kernel: LBO = one 32-channel block (16 KB), SBO = 512 (rows contiguous), layout type 1
    rows = _section("tf32 MN-major, 128B swizzle with 32-byte base, LBO 16384 SBO 512")
    assert len(rows) >= 10
    for m, offs in rows.items():                 # m = channel (M index), k = pixel row inside the 8-row K step
        for k, measured in enumerate(offs):
            assert chunk_addr_sw(k, m // 4) + (m % 4) * 4 == measured, (m, k)


@pytest.mark.skipif(not os.path.exists(RECORD), reason="probe record not present")
def test_other_mn_major_tf32_forms_read_nothing():
    for title in ("tf32 MN-major, no swizzle, LBO 128 SBO 2048", "tf32 MN-major, 128B swizzle, LBO 16384 SBO 1024",
                  "tf32 MN-major, 64B swizzle, LBO 8192 SBO 512", "tf32 MN-major, 32B swizzle, LBO 4096 SBO 256"):
        rows = _section(title)
        assert rows and all(v == 0 for offs in rows.values() for v in offs), title
    # while the K-major form shares the bf16 byte layout: 16-byte chunks 2 KB apart, rows 16 bytes apart
    rows = _section("tf32 K-major, no swizzle, LBO 2048 SBO 128 (forward path)")
    for m, offs in rows.items():
        assert offs == [m * 16 + (k // 4) * 2048 + (k % 4) * 4 for k in range(8)]
