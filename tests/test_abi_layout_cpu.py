"""The ctypes mirror (lvc_b200/_lib.py) and the stub printed in INTEGRATION.md must lay the C structs out exactly
as include/lvc_b200.h does: a tiny C program compiled here with gcc reports sizeof / offsetof of every field."""
import ctypes as C
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STRUCTS = {"lvc_rng": "Rng", "lvc_video": "Video", "lvc_clip_desc": "ClipDesc", "lvc_letterbox_geom": "LetterboxGeom",
           "lvc_embed_desc": "EmbedDesc", "lvc_clip_plan": "ClipPlan"}


def _c_layout(tmp_path):
    from lvc_b200 import _lib

    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "lvc_b200.h"', "int main(void) {"]
    for cname, pyname in STRUCTS.items():
        lines.append(f'  printf("{cname} size %zu\\n", sizeof({cname}));')
        for fname, _ in getattr(_lib, pyname)._fields_:
            lines.append(f'  printf("{cname} {fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ["  return 0;", "}"]
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    table = {}
    for ln in out.splitlines():
        s, f, v = ln.split()
        table[s, f] = int(v)
    return table


@pytest.mark.skipif(shutil.which("gcc") is None, reason="gcc not available")
def test_ctypes_structs_match_the_header(tmp_path):
    from lvc_b200 import _lib

    table = _c_layout(tmp_path)
    for cname, pyname in STRUCTS.items():
        st = getattr(_lib, pyname)
        assert C.sizeof(st) == table[cname, "size"], cname
        for fname, _ in st._fields_:
            assert getattr(st, fname).offset == table[cname, fname], (cname, fname)


@pytest.mark.skipif(shutil.which("gcc") is None, reason="gcc not available")
def test_integration_md_stub_matches_the_header(tmp_path):
    """The struct definitions a maintainer would paste from INTEGRATION.md are executed and compared too."""
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    block = re.search(r"class lvc_rng\(C\.Structure\):.*?\n(?=lib\.lvc_last_error)", text, re.S).group(0)
    ns = {"C": C}
    exec(block, ns)  # noqa: S102 - our own documentation
    table = _c_layout(tmp_path)
    for cname in ("lvc_rng", "lvc_video", "lvc_clip_desc"):
        st = ns[cname]
        assert C.sizeof(st) == table[cname, "size"], cname
        for fname, _ in st._fields_:
            assert getattr(st, fname).offset == table[cname, fname], (cname, fname)
