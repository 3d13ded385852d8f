"""Every script under bench/, tools/, tutorials/ and baseline/ must at least byte-compile (they only run on GPU
boxes, so a syntax error would otherwise surface in the middle of a measurement session)."""
import pathlib
import py_compile

import pytest

ROOT = pathlib.Path(__file__).resolve().parents[1]
SCRIPTS = sorted(p for d in ("bench", "tools", "tutorials", "baseline") for p in (ROOT / d).rglob("*.py")
                 if "_ref" not in p.parts)


@pytest.mark.parametrize("path", SCRIPTS, ids=lambda p: str(p.relative_to(ROOT)))
def test_script_compiles(path, tmp_path):
    py_compile.compile(str(path), cfile=str(tmp_path / "out.pyc"), doraise=True)


def test_entry_points_compile(tmp_path):
    for name in ("bench.py", "__graft_entry__.py"):
        py_compile.compile(str(ROOT / name), cfile=str(tmp_path / (name + "c")), doraise=True)
