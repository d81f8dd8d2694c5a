"""Build libdvsof.so (sm_100a) in-tree with nvcc.  Used by __graft_entry__.build()."""
import os
import subprocess
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / 'csrc'
LIB = Path(os.environ['DVSOF_LIB']) if os.environ.get('DVSOF_LIB') else PKG / 'libdvsof.so'
SOURCES = ['runtime.cu', 'encode.cu', 'transforms.cu', 'loss.cu', 'eval.cu', 'glue.cu']
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '-Xcompiler', '-fPIC', '--expt-relaxed-constexpr']


def _nvcc():
    for cand in (os.environ.get('NVCC'), '/usr/local/cuda/bin/nvcc', 'nvcc'):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    raise RuntimeError('nvcc not found')


def needs_build():
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    deps = [CSRC / s for s in SOURCES] + sorted(CSRC.glob('*.cuh')) + sorted(CSRC.glob('*.h')) + \
        [PKG.parent / 'include' / 'dvsof.h']
    return any(d.stat().st_mtime > t for d in deps)


def build_lib(force=False, verbose=False):
    """Compile every .cu under csrc/ for sm_100a and link libdvsof.so next to this file."""
    if not force and not needs_build():
        return LIB
    nvcc = _nvcc()
    objdir = PKG / 'build'
    objdir.mkdir(exist_ok=True)
    objs = []
    procs = []
    for src in SOURCES:
        obj = objdir / (src[:-3] + '.o')
        cmd = [nvcc, *NVCC_FLAGS, *os.environ.get('DVSOF_NVCC_EXTRA', '').split(), '-c', str(CSRC / src),
               '-o', str(obj)]
        if verbose:
            print(' '.join(cmd))
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
        objs.append(str(obj))
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode:
            raise RuntimeError(f'nvcc failed on {src}:\n{out.decode()}')
    cmd = [nvcc, '-shared', '-o', str(LIB), *objs, '-gencode', 'arch=compute_100a,code=sm_100a']
    subprocess.run(cmd, check=True)
    return LIB


if __name__ == '__main__':
    print(build_lib(force=True, verbose=True))
