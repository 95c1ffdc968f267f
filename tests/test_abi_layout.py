"""The ctypes mirrors of the C ABI (`ops/native.py`, `parallel/dp.py`) against `include/dtf_b200.h`: a small C
program compiled with the host compiler prints sizeof / offsetof of every mirrored struct field, and both sides must
agree.  Guards the hand-maintained Python structs against header drift (CPU-only, no CUDA needed: the header compiles
as plain C when `__CUDACC__` is not defined)."""
import ctypes as C
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _mirrors():
    from distributedtf_b200.ops import native
    from distributedtf_b200.parallel import dp
    return {"DtfBnRef": native.BnRef, "DtfGather": native.Gather, "DtfConvGemm": native.ConvGemm,
            "DtfConvWgrad": native.ConvWgrad, "DtfHead": native.Head, "DtfHparams": native.Hparams,
            "DtfOptSegment": native.OptSegment, "DtfInitSegment": native.InitSegment, "DtfGemm": native.Gemm,
            "DtfAllreduceUpdate": dp.AllreduceUpdate}


@pytest.mark.skipif(shutil.which("gcc") is None and shutil.which("cc") is None, reason="no host C compiler")
def test_ctypes_structs_match_the_header(tmp_path):
    mirrors = _mirrors()
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include <stdint.h>', '#include "dtf_b200.h"', "int main(void) {"]
    for cname, cls in mirrors.items():
        lines.append(f'  printf("{cname} sizeof %zu\\n", sizeof({cname}));')
        for fname, _ in cls._fields_:
            lines.append(f'  printf("{cname} {fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ["  return 0;", "}"]
    src = tmp_path / "abi.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "abi"
    cc = shutil.which("gcc") or shutil.which("cc")
    r = subprocess.run([cc, "-std=c11", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout
    seen = 0
    for line in out.splitlines():
        cname, field, value = line.split()
        cls = mirrors[cname]
        want = C.sizeof(cls) if field == "sizeof" else getattr(cls, field).offset
        assert int(value) == want, f"{cname}.{field}: header {value}, ctypes {want}"
        seen += 1
    assert seen == sum(len(c._fields_) + 1 for c in mirrors.values())
