"""Build libfiode_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m fiode_b200.build            # incremental
    python -m fiode_b200.build --force

The .so sits next to this file so that it travels with a repo snapshot (it is git-ignored).
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / 'csrc'
INCLUDE = HERE.parent / 'include'
OBJ = CSRC / 'build'
LIB = HERE / 'libfiode_b200.so'

NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '-Xcompiler', '-fPIC', '-Xcompiler', '-fvisibility=hidden', '--expt-relaxed-constexpr']


EXTRA = os.environ.get('FIODE_NVCC_EXTRA', '').split()


def _nvcc() -> str:
    for c in (os.environ.get('NVCC'), '/usr/local/cuda/bin/nvcc', 'nvcc'):
        if c and (os.path.sep not in c or os.path.exists(c)):
            return c
    raise RuntimeError('nvcc not found')


def _stale(out: Path, deps) -> bool:
    if not out.exists():
        return True
    m = out.stat().st_mtime
    return any(Path(d).stat().st_mtime > m for d in deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    srcs = sorted(CSRC.glob('*.cu'))
    hdrs = sorted(CSRC.glob('*.cuh')) + sorted(INCLUDE.glob('*.h'))
    OBJ.mkdir(exist_ok=True)
    nvcc = _nvcc()
    jobs = []
    for s in srcs:
        o = OBJ / (s.stem + '.o')
        if force or _stale(o, [s] + hdrs):
            jobs.append((s, o))

    def compile_one(job):
        s, o = job
        cmd = [nvcc] + NVCC_FLAGS + EXTRA + ['-I', str(INCLUDE), '-c', str(s), '-o', str(o)]
        if verbose:
            cmd.insert(1, '-Xptxas=-v')
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f'nvcc failed for {s.name}:\n{r.stdout}\n{r.stderr}')
        if verbose:
            print(r.stderr)
        return o

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(compile_one, jobs))
    objs = [OBJ / (s.stem + '.o') for s in srcs]
    if force or jobs or _stale(LIB, objs):
        cmd = [nvcc, '-shared', '-o', str(LIB)] + [str(o) for o in objs] + ['-lcudart']
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f'link failed:\n{r.stdout}\n{r.stderr}')
    return LIB


if __name__ == '__main__':
    p = build(force='--force' in sys.argv, verbose='-v' in sys.argv)
    print(p)
