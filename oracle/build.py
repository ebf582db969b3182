"""oracle/build.py -- TEST INFRASTRUCTURE ONLY.  Compiles the oracle's plain-C pieces with gcc.

    python -m oracle.build

Output goes to oracle/_build/ (git-ignored as *.so; it still travels to the GPU box with the snapshot).
The reference itself is pure Python (nothing of it compiles), so there is no oracle/_ref recipe.
"""
from __future__ import annotations

import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
OUT = HERE / '_build'
LIB = OUT / 'liboracle_grid.so'


def build(force: bool = False) -> Path:
    src = HERE / 'grid_enum.c'
    OUT.mkdir(exist_ok=True)
    if force or not LIB.exists() or LIB.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(['gcc', '-O2', '-shared', '-fPIC', str(src), '-o', str(LIB)], check=True)
    return LIB


if __name__ == '__main__':
    print(build(force=True))
