"""The ctypes mirrors in spmv_research_b200/capi.py have the same size and field offsets as the C structs of
include/b200spmv.h (checked by compiling a tiny C program with gcc against the header)."""
import ctypes as C
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _c_layout(tmp_path, struct, fields):
    src = tmp_path / ("layout_%s.c" % struct)
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "b200spmv.h"', 'int main(void) {',
             '  printf("sizeof %%zu\\n", sizeof(%s));' % struct]
    for f in fields:
        lines.append('  printf("%s %%zu\\n", offsetof(%s, %s));' % (f, struct, f))
    lines += ['  return 0;', '}']
    src.write_text("\n".join(lines))
    exe = tmp_path / ("layout_%s" % struct)
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)])
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout
    return dict((k, int(v)) for k, v in (ln.split() for ln in out.splitlines()))


def _check(tmp_path, struct, cls):
    names = [n for n, _ in cls._fields_]
    lay = _c_layout(tmp_path, struct, names)
    assert lay["sizeof"] == C.sizeof(cls), (struct, lay["sizeof"], C.sizeof(cls))
    for n in names:
        assert lay[n] == getattr(cls, n).offset, (struct, n, lay[n], getattr(cls, n).offset)


def test_struct_layouts_match_the_header(native, tmp_path):
    from spmv_research_b200 import capi
    _check(tmp_path, "b200spmv_opts", capi.Opts)
    _check(tmp_path, "b200spmv_launch", capi.Launch)
    _check(tmp_path, "b200spmv_pcg_result", capi.PcgResult)
    _check(tmp_path, "b200spmv_power_opts", capi.PowerOpts)


def test_every_opts_field_of_the_header_is_mirrored(native):
    from spmv_research_b200 import capi
    src = open(os.path.join(ROOT, "include", "b200spmv.h")).read()
    body = re.search(r"typedef struct b200spmv_opts \{(.*?)\} b200spmv_opts;", src, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    c_fields = re.findall(r"int32_t\s+(\w+)", body)
    assert c_fields == [n for n, _ in capi.Opts._fields_]
